"""Micro-benchmark of the pooling kernels on one synthetic scene (device-resident): plan build, per-application time."""
import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from geopurify_b200 import synth, ops, _lib as L
import ctypes

n = int(sys.argv[1]) if len(sys.argv) > 1 else 150000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 96
T = int(sys.argv[3]) if len(sys.argv) > 3 else 19
sc = synth.make_scene(n_points=n, n_views=2, feat_dim=16)
vox = torch.as_tensor(sc.voxel_coords.astype(np.int32)).cuda()
m = vox.shape[0]
idx = ops.VoxelIndex(vox)
_, nbr = idx.knn(K, want_orig=False)
emb = torch.nn.functional.normalize(torch.randn(m, 128, device="cuda"), dim=1)
w = ops.affinity_softmax(emb, nbr, 20.0)
f = torch.randn(m, 544, device="cuda")
lib = L.load()

def timeit(fn, reps=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

nb = lib.gp_pool_plan_bytes(m, K)
plan = torch.empty(nb, dtype=torch.uint8, device="cuda")
t_plan = timeit(lambda: L.check(lib.gp_pool_plan(L.ptr(nbr), L.ptr(w), m, K, L.ptr(plan), nb, L.stream())))
tiles = (m + 127) // 128
cnt = plan[: 4 * tiles].view(torch.int32).float()
print(f"M={m} K={K} plan {t_plan:.3f} ms ({nb/1e6:.0f} MB virtual), union per 128-row tile: mean {cnt.mean():.1f} max {cnt.max():.0f}")
t_row = timeit(lambda: ops.pool(f, 512, nbr, w, T, tensor_core=False))
t_tc = timeit(lambda: ops.pool(f, 512, nbr, w, T, tensor_core=True))
print(f"rowwise {t_row:.3f} ms ({t_row / T:.3f}/app)  tensor-core incl. plan {t_tc:.3f} ms ({(t_tc - t_plan) / T:.3f}/app)")
a = ops.pool(f, 512, nbr, w, T, tensor_core=False); b = ops.pool(f, 512, nbr, w, T, tensor_core=True)
print("max diff", (a - b).abs().max().item())
