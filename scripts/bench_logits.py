"""Classification micro-benchmark: tcgen05 GEMM + fused arg-max vs the fp32 SIMT kernel, M = 130k rows, D = 512."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from geopurify_b200 import ops

rows, D = 130_210, 512
feat = torch.randn(rows, D, device="cuda") * 0.05


def timeit(fn, reps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for C in (20, 160, 200, 256):
    text = torch.randn(C, D, device="cuda")
    t_tc = timeit(lambda: ops.logits_argmax(feat, D, text, 100.0, want_zero=False, tensor_core=True))
    t_simt = timeit(lambda: ops.logits_argmax(feat, D, text, 100.0, want_zero=False, tensor_core=False))
    a = ops.logits_argmax(feat, D, text, 100.0, want_zero=False, tensor_core=True)[0]
    b = ops.logits_argmax(feat, D, text, 100.0, want_zero=False, tensor_core=False)[0]
    gb = (rows * D * 4 + C * D * 4 + rows * 4) / 1e9
    print(f"C={C}: tcgen05 {t_tc:.3f} ms ({gb / t_tc * 1e3:.0f} GB/s alg.), fp32 SIMT {t_simt:.3f} ms, "
          f"labels equal: {bool(torch.equal(a, b))}")
