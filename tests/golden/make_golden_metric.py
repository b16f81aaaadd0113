"""Golden vectors for SURVEY.md §8 row A12 (metric part) from the REFERENCE'S OWN `util/util.py`.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_metric.py

`util/util.py` imports matplotlib at module level (absent here; only its plotting helpers use it) and
`intersectionAndUnionGPU` ends with `.cuda()` (no GPU in this container).  Both are environment shims, not algorithm
changes: stub modules stand in for matplotlib and `Tensor.cuda` is made the identity while the reference function
runs; its ignore overwrite and the three `torch.histc` calls execute unmodified.  Only inputs and outputs are stored.
"""
import os
import sys
import types

import numpy as np
import torch

for name in ("matplotlib", "matplotlib.patches", "matplotlib.pyplot"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.path.insert(0, "/root/reference")
from util import util as ref_util  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def run(pred, target, k, ignore):
    cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **kw: self
    try:
        out = torch.as_tensor(pred).clone()
        i, u, t = ref_util.intersectionAndUnionGPU(out, torch.as_tensor(target), k, ignore)
    finally:
        torch.Tensor.cuda = cuda
    return dict(pred=pred, target=target, k=k, ignore=np.array(ignore, dtype=np.int64), pred_after=out.numpy(),
                inter=i.numpy(), union=u.numpy(), tgt=t.numpy())


def main():
    rng = np.random.default_rng(5557)
    cases = {}
    n = 20000
    tgt = rng.integers(0, 20, n)
    tgt[rng.random(n) < 0.05] = 255
    pred = np.where(rng.random(n) < 0.6, np.clip(tgt, 0, 19), rng.integers(0, 20, n))
    cases["scannet20"] = run(pred.astype(np.int64), tgt.astype(np.int64), 20, [255])
    tgt = rng.integers(0, 200, n)
    tgt[rng.random(n) < 0.03] = 255
    tgt[rng.random(n) < 0.03] = 254
    tgt[rng.random(n) < 0.02] = 230          # a label >= K that is NOT an ignore id: dropped by the histc range
    pred = np.where(rng.random(n) < 0.3, np.clip(tgt, 0, 199), rng.integers(0, 200, n))
    cases["scannet200_two_ignores"] = run(pred.astype(np.int64), tgt.astype(np.int64), 200, [255, 254])
    cases["single_point"] = run(np.array([3], dtype=np.int64), np.array([3], dtype=np.int64), 7, [255])
    cases["all_ignored"] = run(rng.integers(0, 5, 50).astype(np.int64), np.full(50, 255, dtype=np.int64), 5, [255])
    for name, c in cases.items():
        np.savez_compressed(os.path.join(HERE, f"metric_{name}.npz"), **c)
        print(name, c["inter"].sum(), c["union"].sum(), c["tgt"].sum())


if __name__ == "__main__":
    main()
