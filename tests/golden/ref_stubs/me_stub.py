"""The subset of MinkowskiEngine 0.5 that `models/affinity_module.py` touches on the inference path, written from the
library's documented semantics (not its code):

* `SparseTensor(features, coordinates)` with unique coordinates keeps the given row order; `.F`, `.C`; `a += b` and
  `a + b` add features of tensors that share a coordinate map;
* `utils.batched_coordinates([c])` prepends the batch index column;
* `MinkowskiConvolution(cin, cout, kernel_size, dimension=3)`: stride 1, no bias, parameter `.kernel` of shape
  `[27, cin, cout]` (`[cin, cout]` for kernel_size 1); out[i] = sum over the occupied offsets of in[j] @ kernel[o],
  offsets enumerated x fastest: o = (dx+1) + 3(dy+1) + 9(dz+1); accumulated offset by offset in float32;
* `MinkowskiBatchNorm(c)` = `.bn = BatchNorm1d(c)` on the feature matrix; `MinkowskiReLU`, `MinkowskiFunctional.relu`.
"""
import types

import numpy as np
import torch
import torch.nn as nn


class SparseTensor:
    def __init__(self, features, coordinates=None, device=None, coordinate_map_key=None, coordinate_manager=None,
                 _nbr=None):
        self.F = features
        self.C = coordinates
        self._nbr = _nbr          # cached 3^3 neighbour table, shared by the tensors of one coordinate map

    @property
    def device(self):
        return self.F.device

    def _like(self, feats):
        return SparseTensor(feats, self.C, _nbr=self._nbr)

    def __add__(self, other):
        return self._like(self.F + other.F)

    def __iadd__(self, other):
        self.F = self.F + other.F
        return self

    def neighbours(self):
        if self._nbr is None:
            c = self.C[:, 1:].cpu().numpy().astype(np.int64) + 1
            assert (self.C[:, 0] == self.C[0, 0]).all(), "single batch only"
            ext = c.max(0) + 2
            key = (c[:, 2] * ext[1] + c[:, 1]) * ext[0] + c[:, 0]
            assert np.unique(key).shape[0] == key.shape[0], "coordinates must be unique"
            order = np.argsort(key, kind="stable")
            skey = key[order]
            m = c.shape[0]
            nbr = np.full((m, 27), -1, dtype=np.int64)
            for o in range(27):
                dx, dy, dz = o % 3 - 1, (o // 3) % 3 - 1, o // 9 - 1
                q = key + (dz * ext[1] + dy) * ext[0] + dx
                pos = np.minimum(np.searchsorted(skey, q), m - 1)
                hit = skey[pos] == q
                nbr[hit, o] = order[pos[hit]]
            self._nbr = torch.from_numpy(nbr)
        return self._nbr


def _batched_coordinates(coords, dtype=torch.int32, device=None):
    out = []
    for b, c in enumerate(coords):
        c = torch.as_tensor(c).to(dtype)
        out.append(torch.cat([torch.full((c.shape[0], 1), b, dtype=dtype), c.cpu()], 1))
    return torch.cat(out, 0)


utils = types.SimpleNamespace(batched_coordinates=_batched_coordinates)


class MinkowskiConvolution(nn.Module):
    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False, dimension=None):
        super().__init__()
        assert dimension == 3 and stride == 1 and dilation == 1 and not bias and kernel_size in (1, 3)
        self.kernel_size = kernel_size
        shape = (in_channels, out_channels) if kernel_size == 1 else (27, in_channels, out_channels)
        self.kernel = nn.Parameter(torch.zeros(shape))

    def forward(self, x):
        if self.kernel_size == 1:
            return x._like(x.F @ self.kernel)
        nbr = x.neighbours()
        out = torch.zeros((x.F.shape[0], self.kernel.shape[2]), dtype=x.F.dtype)
        for o in range(27):
            rows = torch.where(nbr[:, o] >= 0)[0]
            if rows.numel():
                out.index_add_(0, rows, x.F[nbr[rows, o]] @ self.kernel[o])
        return x._like(out)


class MinkowskiBatchNorm(nn.Module):
    def __init__(self, num_features, eps=1e-5, momentum=0.1):
        super().__init__()
        self.bn = nn.BatchNorm1d(num_features, eps=eps, momentum=momentum)

    def forward(self, x):
        return x._like(self.bn(x.F))


class MinkowskiReLU(nn.Module):
    def forward(self, x):
        return x._like(torch.relu(x.F))


MinkowskiFunctional = types.ModuleType("MinkowskiEngine.MinkowskiFunctional")
MinkowskiFunctional.relu = lambda x: x._like(torch.relu(x.F))
