"""A5 micro-benchmark: consensus top-3 lifting (gp_lift_consensus + scene-level fill) on the N=150k scene, against the
A4 mean lifting from the same compact 2-D form."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from geopurify_b200 import synth, ops
from geopurify_b200.affinity_module import scene_inputs_from_synth

sc = synth.make_scene(n_points=150_000, n_views=50, image_dim=(320, 240), feat_dim=512, n_classes=20, seed=5557)
sc.tables[:, 0, :] = 0.0                      # mask 0 of every view = "no mask": exercises the per-view fills
s = scene_inputs_from_synth(sc, feat_dtype=torch.bfloat16)
W, H = sc.image_dim
text = torch.as_tensor(sc.text).float().cuda()
pix, counts = ops.project_views(s.coords, s.w2c, s.intr, s.depth, H, W, s.cut, s.vis_thres)
keep = ops.accept_views(counts, 400)


def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def consensus():
    out, seen = ops.lift_consensus(pix, keep, s.idmaps, s.tables, text, 100.0, s.coords_f32)
    nn = ops.nn1_fill(s.coords_f32, seen)
    ops.copy_rows(out, 512, nn)
    return out


def mean():
    out, seen = ops.lift_mean_idmap(s.coords, s.w2c, s.intr, s.depth, H, W, s.cut, s.vis_thres, keep, s.idmaps, s.tables)
    nn = ops.nn1_fill(s.coords_f32, seen)
    ops.copy_rows(out, 512, nn)
    return out


zero_pairs = int(((pix >= 0) & keep.bool()[:, None] & (s.idmaps.view(50, -1).gather(1, pix.clamp(min=0).long()) == 0)).sum())
print(f"visible pairs {int((pix >= 0).sum())}, under a zero mask {zero_pairs}")
print(f"consensus lifting + fill {timeit(consensus):.3f} ms;  mean lifting + fill {timeit(mean):.3f} ms")
