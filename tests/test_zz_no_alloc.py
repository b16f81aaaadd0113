"""gp_scatter_mean_static: the allocation-free variant of A6 (scratch in a static __device__ array inside the library).

Named test_zz_* so that pytest collects it LAST (`pytest --collect-only -q tests` shows the order):
a failure here must not hide another test.  It has no counterpart in the reference,
so its check is the existing entry points gp_scatter_mean_prepare + gp_scatter_mean called
through the same C ABI: the same kernels run on the same layout, hence the tolerance is the one the scatter-mean
parity tests use -- bit-exact (torch.equal).  Two calls in a row with different sizes share the static array; the
second must not be disturbed by what the first left behind.  The test needs no fixture from the other files and
passes alone in a fresh process as well as after the whole suite.
"""
from ctypes import c_int64, c_size_t

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_N = 262144   # GP_SCATTER_STATIC_MAX_N / _MAX_M in include/geopurify_b200.h
MAX_M = 262144


def _case(n, m, D, stride, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    p2v = torch.randint(0, m, (n,), device="cuda", generator=g, dtype=torch.int64).to(torch.int32)
    if m > 8:                       # empty destination rows at both ends and in the middle
        p2v[(p2v == 0) | (p2v == m - 1) | (p2v == m // 2)] = 1
    src = torch.randn(n, stride, device="cuda", generator=g, dtype=torch.float32)
    return p2v.contiguous(), src


def _existing(L, lib, p2v, src, n, m, D, out_stride, col0):
    """The existing path through the C ABI: caller-allocated workspace (torch) + prepare + scatter."""
    nb = int(lib.gp_scatter_mean_workspace_bytes(c_int64(n), c_int64(m)))
    ws = torch.empty(max(nb, 256), dtype=torch.uint8, device="cuda")
    out = torch.full((m, out_stride), -7.0, device="cuda", dtype=torch.float32)
    L.check(lib.gp_scatter_mean_prepare(L.ptr(p2v, torch.int32), n, m, L.ptr(ws), c_size_t(nb), L.stream()))
    L.check(lib.gp_scatter_mean(L.ptr(src, torch.float32), src.stride(0), D, L.ptr(ws), n, m, L.ptr(out), out_stride,
                                col0, L.stream()))
    return out


def _static(L, lib, p2v, src, n, m, D, out, col0):
    L.check(lib.gp_scatter_mean_static(L.ptr(p2v, torch.int32), n, m, L.ptr(src, torch.float32), src.stride(0), D,
                                       L.ptr(out, torch.float32), out.stride(0), col0, L.stream()))


def test_scatter_mean_static_matches_workspace_path_on_two_calls_in_a_row():
    from geopurify_b200 import _lib as L
    L.require_cuda()
    lib = L.load()
    # call 1: a full-size scene's shape (518 feature columns into a 544-wide row at column 0, as the path does);
    # call 2: smaller, different data and geometry, scalar (non-float4) code path, lands on call 1's leftovers
    cases = [dict(n=150_000, m=130_000, D=518, stride=520, out_stride=544, col0=0, seed=11),
             dict(n=40_001, m=9_173, D=7, stride=9, out_stride=13, col0=3, seed=12)]
    inputs, want = [], []
    for c in cases:
        p2v, src = _case(c["n"], c["m"], c["D"], c["stride"], c["seed"])
        inputs.append((p2v, src))
        want.append(_existing(L, lib, p2v, src, c["n"], c["m"], c["D"], c["out_stride"], c["col0"]))
    outs = [torch.full((c["m"], c["out_stride"]), -7.0, device="cuda", dtype=torch.float32) for c in cases]
    torch.cuda.synchronize()

    # the two calls back to back on one stream, nothing in between
    for c, (p2v, src), out in zip(cases, inputs, outs):
        _static(L, lib, p2v, src, c["n"], c["m"], c["D"], out, c["col0"])
    torch.cuda.synchronize()
    for c, w, out in zip(cases, want, outs):
        assert torch.equal(out, w), f"static scratch result differs from the workspace path for {c}"
        # sanity of the comparison itself: columns outside [col0, col0 + D) untouched, inside written
        assert bool((out[:, c["col0"] + c["D"]:] == -7.0).all()) and bool((out[:, :c["col0"]] == -7.0).all())
    # an independent check of call 2 against torch (fp32 sums of <= ~10 rows: tight tolerance, order-free)
    c, (p2v, src), out = cases[1], inputs[1], outs[1]
    ref = torch.zeros(c["m"], c["D"], device="cuda", dtype=torch.float64)
    ref.index_add_(0, p2v.long(), src[:, :c["D"]].double())
    cnt = torch.bincount(p2v.long(), minlength=c["m"]).clamp(min=1).double()[:, None]
    assert (out[:, c["col0"]:c["col0"] + c["D"]].double() - ref / cnt).abs().max().item() < 1e-5

    # no allocation: repeat call 2 and watch the driver's free memory and torch's allocator
    free0, _ = torch.cuda.mem_get_info()
    stats0 = torch.cuda.memory_stats()
    _static(L, lib, p2v, src, c["n"], c["m"], c["D"], out, c["col0"])
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    stats1 = torch.cuda.memory_stats()
    assert free1 == free0, "device memory was allocated or released during gp_scatter_mean_static"
    assert stats1.get("allocation.all.allocated", 0) == stats0.get("allocation.all.allocated", 0)
    assert stats1.get("segment.all.allocated", 0) == stats0.get("segment.all.allocated", 0)
    assert torch.equal(out, want[1])


def test_scatter_mean_static_limits():
    from geopurify_b200 import _lib as L
    L.require_cuda()
    lib = L.load()
    # at the limits it runs (the layout for MAX_N, MAX_M fits the 16 MiB array) and agrees with the workspace path
    D = 4
    p2v, src = _case(MAX_N, MAX_M, D, D, 13)
    want = _existing(L, lib, p2v, src, MAX_N, MAX_M, D, D, 0)
    out = torch.empty(MAX_M, D, device="cuda", dtype=torch.float32)
    _static(L, lib, p2v, src, MAX_N, MAX_M, D, out, 0)
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    # beyond them it refuses with a clear message before touching any buffer
    for n, m in ((MAX_N + 1, 16), (16, MAX_M + 1)):
        with pytest.raises(L.GeoPurifyError, match="beyond the static scratch limits"):
            _static(L, lib, p2v, src, n, m, D, out, 0)
    # empty inputs: n = 0 gives all-zero rows (count clamps to 1), m = 0 writes nothing
    out.fill_(5.0)
    _static(L, lib, p2v, src, 0, 32, D, out, 0)
    torch.cuda.synchronize()
    assert bool((out[:32] == 0).all()) and bool((out[32:] == 5.0).all())
    _static(L, lib, p2v, src, 0, 0, D, out, 0)
    torch.cuda.synchronize()
