"""Check (not a pytest test: needs two GPUs) that gp_scatter_mean_static resolves its static scratch per device:
run under `gpurun --gpus 2`; prints `equal True` for devices 1, 0, 1 of ONE process (verified on 2 x B200)."""
import sys, os, torch
sys.path.insert(0, os.getcwd())
from ctypes import c_int64, c_size_t
from geopurify_b200 import _lib as L
L.require_cuda(); lib = L.load()
for dev in (1, 0, 1):
    torch.cuda.set_device(dev)
    n, m, D = 50000, 20000, 32
    g = torch.Generator(device=f"cuda:{dev}").manual_seed(dev + 7)
    p2v = torch.randint(0, m, (n,), device=f"cuda:{dev}", generator=g).to(torch.int32)
    src = torch.randn(n, D, device=f"cuda:{dev}", generator=g)
    nb = int(lib.gp_scatter_mean_workspace_bytes(c_int64(n), c_int64(m)))
    ws = torch.empty(nb, dtype=torch.uint8, device=f"cuda:{dev}")
    want = torch.empty(m, D, device=f"cuda:{dev}"); out = torch.empty_like(want)
    L.check(lib.gp_scatter_mean_prepare(L.ptr(p2v), n, m, L.ptr(ws), c_size_t(nb), L.stream()))
    L.check(lib.gp_scatter_mean(L.ptr(src), D, D, L.ptr(ws), n, m, L.ptr(want), D, 0, L.stream()))
    L.check(lib.gp_scatter_mean_static(L.ptr(p2v), n, m, L.ptr(src), D, D, L.ptr(out), D, 0, L.stream()))
    torch.cuda.synchronize()
    print("device", dev, "equal", torch.equal(out, want))
