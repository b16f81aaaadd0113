"""Per-scene timing of the hole fill (1-NN over seen points) on the bench's scene pool."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from geopurify_b200 import synth, ops
from geopurify_b200.affinity_module import scene_inputs_from_synth

rng = np.random.default_rng(5557)
sizes = [150_000] + [synth.scannet_val_size(rng) for _ in range(3)]
for i, n in enumerate(sizes):
    sc = synth.make_scene(n_points=n, n_views=50, image_dim=(320, 240), feat_dim=16, n_classes=20, seed=5557 + i)
    s = scene_inputs_from_synth(sc, feat_dtype=torch.float32)
    W, H = sc.image_dim
    pix, counts = ops.project_views(s.coords, s.w2c, s.intr, s.depth, H, W, s.cut, s.vis_thres)
    keep = ops.accept_views(counts, 400)
    seen = ((pix >= 0) & keep.bool()[:, None]).any(0).to(torch.uint8).contiguous()
    for _ in range(2):
        nn = ops.nn1_fill(s.coords_f32, seen, 0.05)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(5):
        nn = ops.nn1_fill(s.coords_f32, seen, 0.05)
    e1.record(); torch.cuda.synchronize()
    uns = int((seen == 0).sum())
    d = (s.coords_f32 - s.coords_f32[nn.long()]).norm(dim=1)[seen == 0]
    q = torch.quantile(d, torch.tensor([0.5, 0.9, 0.99, 1.0], device="cuda")).tolist() if uns else []
    print(f"N={n} kept views {int(keep.sum())} unseen {uns} ({100*uns/n:.1f}%) fill {e0.elapsed_time(e1)/5:.3f} ms  nn dist quantiles 50/90/99/100: {q}")
