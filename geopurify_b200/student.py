"""Student affinity network: host-side mirror of the reference's `AffinityPredictor`.

Same constructor, `forward`, `get_param_groups` and state-dict keys as reference
models/affinity_module.py:33-85 (MinkowskiEngine naming: `input_layer.0.kernel`,
`input_layer.1.bn.*`, `res_blocks.N.conv{1,2}.kernel`, `res_blocks.N.norm{1,2}.bn.*`,
`output_layer.kernel`), so the released checkpoint loads with `load_state_dict`.
The forward pass runs in libgeopurify_b200.so (gp_student_forward); inference only.
"""
from __future__ import annotations

import ctypes
import math

import numpy as np
import torch
import torch.nn as nn

from . import _lib as L
from . import ops


class SparseTensor:
    """Minimal stand-in for ME.SparseTensor: `.F` features [M,C], `.C` coordinates [M,4] (batch, x, y, z)."""

    def __init__(self, features, coordinates, device=None, nbr27=None, index=None):
        self.F = features
        self.C = coordinates
        self.nbr27 = nbr27      # optional precomputed kernel map (sorted space) when F is in sorted order
        self.index = index


class SparseConvolution(nn.Module):
    """ME.MinkowskiConvolution(in, out, kernel_size, dimension=3), stride 1, no bias: parameter `kernel`."""

    def __init__(self, in_channels, out_channels, kernel_size=3, dimension=3):
        super().__init__()
        assert dimension == 3 and kernel_size in (1, 3)
        self.in_channels, self.out_channels, self.kernel_size = in_channels, out_channels, kernel_size
        kv = kernel_size ** 3
        shape = (kv, in_channels, out_channels) if kv > 1 else (in_channels, out_channels)
        self.kernel = nn.Parameter(torch.empty(shape))
        nn.init.normal_(self.kernel, 0.0, math.sqrt(2.0 / (kv * in_channels)))


class SparseBatchNorm(nn.Module):
    """ME.MinkowskiBatchNorm: wraps BatchNorm1d under the attribute name `bn`."""

    def __init__(self, channels, eps=1e-5):
        super().__init__()
        self.bn = nn.BatchNorm1d(channels, eps=eps)


class SparseReLU(nn.Module):
    pass


class MinkowskiResBlock(nn.Module):
    def __init__(self, channels):
        super().__init__()
        self.conv1 = SparseConvolution(channels, channels, kernel_size=3, dimension=3)
        self.norm1 = SparseBatchNorm(channels)
        self.conv2 = SparseConvolution(channels, channels, kernel_size=3, dimension=3)
        self.norm2 = SparseBatchNorm(channels)


def _pad32(c):
    return (c + 31) // 32 * 32


class AffinityPredictor(nn.Module):
    def __init__(self, input_dim, embed_dim=128, hidden_dim=256):
        super().__init__()
        self.input_dim, self.embed_dim, self.hidden_dim = input_dim, embed_dim, hidden_dim
        self.input_layer = nn.Sequential(
            SparseConvolution(input_dim, hidden_dim, kernel_size=3, dimension=3),
            SparseBatchNorm(hidden_dim),
            SparseReLU(),
        )
        self.res_blocks = nn.Sequential(*[MinkowskiResBlock(hidden_dim) for _ in range(4)])
        self.output_layer = SparseConvolution(hidden_dim, embed_dim, kernel_size=1, dimension=3)
        self._packed = None
        self._packed_tc = None
        self.precision = "bf16x3"   # "bf16x3": tcgen05 tensor cores, fp32-faithful split; "fp32": SIMT verification path
        self.eval()

    # -- reference API -------------------------------------------------------------------------
    def get_param_groups(self):
        return {"input": list(self.input_layer.parameters()), "middle": list(self.res_blocks.parameters()),
                "output": list(self.output_layer.parameters())}

    def load_state_dict(self, state_dict, strict=True, **kw):
        self._packed = None
        self._packed_tc = None
        return super().load_state_dict(state_dict, strict=strict, **kw)

    @property
    def input_stride(self):
        """Row stride (floats) the packed first layer expects: input_dim padded to a multiple of 32."""
        return _pad32(self.input_dim)

    # -- packing -------------------------------------------------------------------------------
    def _layers(self):
        out = [(self.input_layer[0], self.input_layer[1], True, -1)]
        for b, blk in enumerate(self.res_blocks):
            out.append((blk.conv1, blk.norm1, True, -1))
            out.append((blk.conv2, blk.norm2, True, len(out) - 2))  # residual from layer l-2
        out.append((self.output_layer, None, False, -1))
        return out

    def pack(self, device):
        """Flatten weights (+ folded BN scale/shift) into one device buffer and fill the C descriptor."""
        if self._packed is not None and self._packed[1].device == device:
            return self._packed
        desc = L.StudentDesc()
        chunks, off = [], 0
        layers = self._layers()
        desc.n_layers = len(layers)
        for l, (conv, norm, relu, res) in enumerate(layers):
            w = conv.kernel.detach().to(torch.float32).cpu()
            kv = conv.kernel_size ** 3
            w = w.reshape(kv, conv.in_channels, conv.out_channels)
            cin_pad = _pad32(conv.in_channels)
            if cin_pad != conv.in_channels:
                w = torch.cat([w, torch.zeros(kv, cin_pad - conv.in_channels, conv.out_channels)], 1)
            desc.kv[l], desc.cin[l], desc.cout[l] = kv, cin_pad, conv.out_channels
            desc.relu[l] = int(relu)
            desc.residual_from[l] = l - 2 if res >= 0 else -1
            desc.w_off[l] = off
            chunks.append(w.reshape(-1))
            off += w.numel()
            if norm is not None:
                bn = norm.bn
                var = bn.running_var.detach().double().cpu()
                scale = bn.weight.detach().double().cpu() / torch.sqrt(var + bn.eps)
                shift = bn.bias.detach().double().cpu() - bn.running_mean.detach().double().cpu() * scale
                desc.bn[l] = 1
                desc.bn_off[l] = off
                chunks += [scale.float(), shift.float()]
                off += 2 * conv.out_channels
            else:
                desc.bn[l] = 0
                desc.bn_off[l] = 0
            pad = (-off) % 4   # keep every block 16-byte aligned
            if pad:
                chunks.append(torch.zeros(pad))
                off += pad
        flat = torch.cat(chunks).contiguous().to(device)
        self._packed = (desc, flat)
        return self._packed

    # -- forward -------------------------------------------------------------------------------
    @torch.no_grad()
    def forward_sorted(self, x, nbr27, normalize=True):
        """x [M, >=input_stride] fp32 rows in the voxel index's sorted order (pad columns zero), nbr27 [M,27] int32."""
        L.require_cuda()
        lib = L.load()
        desc, flat = self.pack(x.device)
        m = x.shape[0]
        if x.stride(0) < self.input_stride:
            raise L.GeoPurifyError(f"student input rows need stride >= {self.input_stride} (zero padded)")
        embed = ops._new((m, self.embed_dim), torch.float32, "embed")
        if self.precision == "bf16x3":
            if self._packed_tc is None or self._packed_tc.device != x.device:
                nbp = lib.gp_student_tc_pack_bytes(ctypes.byref(desc))
                self._packed_tc = torch.empty(nbp, dtype=torch.uint8, device=x.device)
                L.check(lib.gp_student_tc_pack(ctypes.byref(desc), L.ptr(flat), L.ptr(self._packed_tc), L.stream()))
                L.count(desc.n_layers)
            nb = lib.gp_student_tc_workspace_bytes(ctypes.byref(desc), m)
            if nb == 0:
                raise L.GeoPurifyError("student: layer structure not supported by the tensor-core path")
            ws = ops._ws(nb, "student_tc")
            L.check(lib.gp_student_forward_tc(ctypes.byref(desc), L.ptr(flat), L.ptr(self._packed_tc), L.ptr(x),
                                              x.stride(0), L.ptr(nbr27), m, L.ptr(embed), int(normalize), L.ptr(ws),
                                              nb, L.stream()))
            nslab = (m + 131071) // 131072
            L.count(6 + 2 * nslab * (desc.n_layers - 1) + 2)
            return embed
        nb = lib.gp_student_workspace_bytes(ctypes.byref(desc), m)
        ws = ops._ws(nb, "student")
        L.check(lib.gp_student_forward(ctypes.byref(desc), L.ptr(flat), L.ptr(x), x.stride(0), L.ptr(nbr27), m,
                                       L.ptr(embed), int(normalize), L.ptr(ws), nb, L.stream()))
        L.count(3 + 27 * 8 + 1 + 9 + 9 + 1)
        return embed

    @torch.no_grad()
    def forward(self, x_s: SparseTensor) -> SparseTensor:
        """Reference signature (models/affinity_module.py:68-72): rows stay in the caller's order."""
        L.require_cuda()
        feats = x_s.F.float().cuda()
        coords = x_s.C[:, -3:].to(torch.int32).cuda().contiguous()
        coords = coords - coords.min(0).values
        index = ops.VoxelIndex(coords)
        nbr27 = index.kernel_map()
        xs = torch.zeros((feats.shape[0], self.input_stride), dtype=torch.float32, device=feats.device)
        xs[:, :feats.shape[1]] = feats[index.order.long()]
        emb = self.forward_sorted(xs, nbr27, normalize=False)
        return SparseTensor(emb[index.rank.long()], x_s.C)


def load_numpy_state(model, state):
    """Load a {name: ndarray} state (geopurify_b200.synth.make_student_state) into an AffinityPredictor."""
    sd = {k: torch.as_tensor(np.asarray(v)) for k, v in state.items()}
    own = model.state_dict()
    for k in own:
        if k.endswith("num_batches_tracked") and k not in sd:
            sd[k] = own[k]
    model.load_state_dict(sd)
    return model
