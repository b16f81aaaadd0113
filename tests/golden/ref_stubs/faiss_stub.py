"""faiss.IndexFlatL2 by its definition: exact squared-L2 search, results in ascending distance.  The coordinates on
this path are small integers stored as float32, so every distance is an exact integer; equal distances are returned
lowest index first (the convention the oracle and the CUDA kernel document for ties)."""
import numpy as np


class IndexFlatL2:
    def __init__(self, d):
        self.d = d
        self.x = np.zeros((0, d), dtype=np.float32)

    def add(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        assert x.shape[1] == self.d
        self.x = np.concatenate([self.x, x], 0)

    def search(self, q, k):
        q = np.ascontiguousarray(q, dtype=np.float32)
        base = self.x.astype(np.float64)
        n = base.shape[0]
        dist = np.empty((q.shape[0], k), dtype=np.float32)
        idx = np.full((q.shape[0], k), -1, dtype=np.int64)
        for s in range(0, q.shape[0], 1024):
            qq = q[s:s + 1024].astype(np.float64)
            d2 = ((qq[:, None, :] - base[None, :, :]) ** 2).sum(-1)
            order = np.argsort(d2, axis=1, kind="stable")[:, :k]        # stable: ties keep the lower index first
            kk = min(k, n)
            idx[s:s + 1024, :kk] = order[:, :kk]
            dist[s:s + 1024, :kk] = np.take_along_axis(d2, order[:, :kk], 1)
        return dist, idx
