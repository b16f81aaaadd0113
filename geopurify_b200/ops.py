"""Thin torch-tensor wrappers over the C ABI: allocate outputs / workspaces, pass pointers, raise on error.

Every function enqueues on the current CUDA stream and returns device tensors.  Like the reference (one Python
thread per process, `torch.no_grad()`), the wrappers are not thread-safe: workspaces come from a per-stream grow-only
arena (see `_ws` / `scratch`), so use one host thread per CUDA stream.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np
import torch

from . import _lib as L


def _dev():
    return torch.device("cuda", torch.cuda.current_device())


# Guard bands around every workspace (GEOPURIFY_B200_GUARD=1, or guard_workspaces(True)): compute-sanitizer is closed on
# the build pool, so out-of-bounds writes of the kernels' internal indexing (student staging, pool plan, voxel index,
# sort scratch) are looked for with 4 KB canaries on both sides of each workspace; check_guards() verifies them.
_GUARD = os.environ.get("GEOPURIFY_B200_GUARD", "0") == "1"
_GUARD_BYTES, _GUARD_BYTE = 4096, 0xA5
_guarded = []


# Scene-internal temporaries (pixel table, lifted rows, voxel rows, kernel map, embeddings, neighbour lists, weights,
# index tables) can come from the same grow-only arena while `scratch()` is active: run_scene / purify switch it on when
# the caller did not ask to keep intermediates, so an evaluation list of differently sized scenes issues no cudaMalloc
# after its largest scene (the caching allocator splits and re-grows its blocks otherwise: ~1.3 cudaMalloc per scene on
# the 312-scene list).  Tensors handed back to the caller (pooled features, labels, seen flags, predictions, histograms)
# are always fresh allocations.
_SCRATCH = False
_ITEM = {torch.uint8: 1, torch.int32: 4, torch.float32: 4, torch.int64: 8, torch.float64: 8, torch.bfloat16: 2}


class scratch:
    def __init__(self, on=True):
        self.on = bool(on)

    def __enter__(self):
        global _SCRATCH
        self.prev = _SCRATCH
        _SCRATCH = self.on or self.prev
        return self

    def __exit__(self, *exc):
        global _SCRATCH
        _SCRATCH = self.prev
        return False


def _new(shape, dtype, tag):
    """Output tensor of a wrapper: a fresh allocation, or (inside `scratch()`) a view of the arena buffer `tag`."""
    if not _SCRATCH or _GUARD:
        return torch.empty(shape, dtype=dtype, device=_dev())
    n = 1
    for d in shape:
        n *= int(d)
    item = _ITEM[dtype]
    return _ws(max(n * item, 256), "out:" + tag)[: n * item].view(dtype).view(shape)


def guard_workspaces(on=True):
    global _GUARD
    _GUARD = bool(on)
    _guarded.clear()


# Call-scoped workspaces come from a grow-only arena, one buffer per (device, stream, call site): scenes of different
# sizes would otherwise ask the caching allocator for a different multi-GB block each (the student's worst-case pair
# staging alone is 54 KB per voxel), which ends in cudaMalloc / cudaFree stalls in the middle of an evaluation list.
# Reuse is safe because a workspace is only touched by kernels the call enqueues on its stream, and the next call with
# the same tag on that stream is ordered behind them.  Workspaces that outlive the call (VoxelIndex, ScatterPlan) are
# plain allocations.
_arena = {}


def _ws(nbytes, tag=None):
    n = max(int(nbytes), 256)
    if not _GUARD:
        if tag is None:
            return torch.empty(n, dtype=torch.uint8, device=_dev())
        key = (torch.cuda.current_device(), torch.cuda.current_stream().cuda_stream, tag)
        buf = _arena.get(key)
        if buf is None or buf.numel() < n:
            _arena.pop(key, None)
            buf = None
            buf = torch.empty(int(n * 1.25) + 4096, dtype=torch.uint8, device=_dev())
            _arena[key] = buf
        return buf[:n]
    n = (n + 255) // 256 * 256
    raw = torch.empty(n + 2 * _GUARD_BYTES, dtype=torch.uint8, device=_dev())
    raw[:_GUARD_BYTES] = _GUARD_BYTE
    raw[_GUARD_BYTES + n:] = _GUARD_BYTE
    _guarded.append((raw, n))
    return raw[_GUARD_BYTES:_GUARD_BYTES + n]      # 4096-byte offset keeps the 256-byte alignment the C ABI asks for


def check_guards():
    """Verify the canaries of every workspace handed out since guard_workspaces(True) (synchronises).
    Returns the number of workspaces checked; raises GeoPurifyError on the first damaged band."""
    torch.cuda.synchronize()
    for i, (raw, n) in enumerate(_guarded):
        lo, hi = raw[:_GUARD_BYTES], raw[_GUARD_BYTES + n:]
        if not bool((lo == _GUARD_BYTE).all()) or not bool((hi == _GUARD_BYTE).all()):
            side = "below" if not bool((lo == _GUARD_BYTE).all()) else "above"
            raise L.GeoPurifyError(f"workspace #{i} ({n} bytes): guard band {side} it was overwritten")
    k = len(_guarded)
    _guarded.clear()
    return k


def cameras_to_device(w2c, intr):
    """w2c [V,4,4] (any float dtype, true world-to-camera), intr [V,4] or [4] (fx,fy,cx,cy) -> float64 device tensors."""
    w = torch.as_tensor(np.asarray(w2c, dtype=np.float64).reshape(-1, 16)).to(_dev())
    k = np.asarray(intr, dtype=np.float64)
    if k.ndim == 1:
        k = np.repeat(k[None, :], w.shape[0], 0)
    return w.contiguous(), torch.as_tensor(np.ascontiguousarray(k)).to(_dev())


def project_views(coords, w2c, intr, depth, H, W, cut, vis_thres):
    """A1: returns pix [V,N] int32 (row*W+col or -1) and counts [V] int32."""
    lib = L.load()
    n, V = coords.shape[0], w2c.shape[0]
    pix = _new((V, n), torch.int32, "pix")
    counts = _new((V,), torch.int32, "counts")
    dd = 0 if depth is None else L.dtype_code(depth)
    L.check(lib.gp_project_views(L.ptr(coords), L.dtype_code(coords), n, L.ptr(w2c), L.ptr(intr), L.ptr(depth), dd, V,
                                 H, W, cut, float(vis_thres), L.ptr(pix), L.ptr(counts), L.stream()))
    L.count(1)
    return pix, counts


def pix_to_mapping(pix, W):
    lib = L.load()
    out = torch.empty(pix.shape + (3,), dtype=torch.int64, device=_dev())
    L.check(lib.gp_pix_to_mapping(L.ptr(pix), pix.numel(), W, L.ptr(out), L.stream()))
    L.count(1)
    return out


def splat_depth(coords, w2c, intr, H, W, cut):
    lib = L.load()
    V = w2c.shape[0]
    depth = torch.empty((V, H, W), dtype=torch.float64, device=_dev())
    L.check(lib.gp_splat_depth(L.ptr(coords), L.dtype_code(coords), coords.shape[0], L.ptr(w2c), L.ptr(intr), V, H, W,
                               cut, L.ptr(depth), L.stream()))
    L.count(2)
    return depth


def accept_views(counts, min_visible=400, max_visible=2**31 - 1):
    lib = L.load()
    keep = _new((counts.shape[0],), torch.uint8, "keep")
    L.check(lib.gp_accept_views(L.ptr(counts), counts.shape[0], int(min_visible), int(max_visible), L.ptr(keep),
                                L.stream()))
    L.count(1)
    return keep


def mask_partition(mask_logits, cls_logits, score_thresh=0.0):
    """F3: [V,Q,H,W] mask logits + [V,Q,C+1] class logits -> (idmap [V,H,W] uint8 with Q = no mask, scores [V,Q])."""
    lib = L.load()
    V, Q, H, W = mask_logits.shape
    idmap = torch.empty((V, H, W), dtype=torch.uint8, device=_dev())
    scores = torch.empty((V, Q), dtype=torch.float32, device=_dev())
    L.check(lib.gp_mask_partition(L.ptr(mask_logits), L.ptr(cls_logits), V, Q, cls_logits.shape[-1], H, W,
                                  float(score_thresh), L.ptr(idmap), L.ptr(scores), L.stream()))
    L.count(2)
    return idmap, scores


def mask_table(embeddings):
    """[V,Q,D] query embeddings -> [V,Q+1,D] table whose last row (index Q = "no mask") is zero."""
    V, Q, D = embeddings.shape
    return torch.cat([embeddings.float(), torch.zeros((V, 1, D), dtype=torch.float32, device=embeddings.device)], 1)


def visible_csr(pix, keep, W, want_pixels=True):
    """F2: CSR of the visible pairs.  Returns (offsets [V+1] int32 device, point_idx [P], view_idx [P], rows [P],
    cols [P]) -- one host read for P.  rows / cols are the reference's x_label / y_label."""
    lib = L.load()
    V, n = pix.shape
    cap = max(V * n, 1)
    off = torch.empty((V + 1,), dtype=torch.int32, device=_dev())
    pidx = torch.empty((cap,), dtype=torch.int32, device=_dev())
    vidx = torch.empty((cap,), dtype=torch.int32, device=_dev())
    rows = torch.empty((cap,), dtype=torch.int32, device=_dev()) if want_pixels else None
    cols = torch.empty((cap,), dtype=torch.int32, device=_dev()) if want_pixels else None
    nb = lib.gp_visible_csr_workspace_bytes(n, V)
    ws = _ws(nb, "visible_csr")
    L.check(lib.gp_visible_csr(L.ptr(pix), L.ptr(keep), n, V, W, L.ptr(off), L.ptr(pidx), L.ptr(vidx), L.ptr(rows),
                               L.ptr(cols), L.ptr(ws), nb, L.stream()))
    L.count(4)
    P = int(off[V].item())
    return off, pidx[:P], vidx[:P], (rows[:P] if want_pixels else None), (cols[:P] if want_pixels else None)


def lift_mean_fused(coords, w2c, intr, depth, H, W, cut, vis_thres, keep, feat, out=None, out_stride=None):
    """A1+A4 fused.  feat [V,H,W,D] fp32/bf16 channels-last.  Returns (out [N,stride] fp32, seen [N] uint8)."""
    lib = L.load()
    n, V, D = coords.shape[0], w2c.shape[0], feat.shape[-1]
    if out is None:
        out_stride = out_stride or D
        out = _new((n, out_stride), torch.float32, "lifted")
    seen = torch.empty((n,), dtype=torch.uint8, device=_dev())
    dd = 0 if depth is None else L.dtype_code(depth)
    L.check(lib.gp_lift_mean_fused(L.ptr(coords), L.dtype_code(coords), n, L.ptr(w2c), L.ptr(intr), L.ptr(depth), dd,
                                   V, H, W, cut, float(vis_thres), L.ptr(keep), L.ptr(feat), L.dtype_code(feat), D,
                                   L.ptr(out), out.stride(0), L.ptr(seen), L.stream()))
    L.count(1)
    return out, seen


def lift_mean_idmap(coords, w2c, intr, depth, H, W, cut, vis_thres, keep, idmap_u8, table, out=None, out_stride=None):
    """A1+A4 fused on the compact 2-D form (idmap [V,H,W] uint8, table [V,Q,D] fp32/bf16): no dense map is built.
    Returns (out [N,stride] fp32, seen [N] uint8)."""
    lib = L.load()
    n, V = coords.shape[0], w2c.shape[0]
    Q, D = table.shape[1], table.shape[2]
    if out is None:
        out_stride = out_stride or D
        out = _new((n, out_stride), torch.float32, "lifted")
    seen = torch.empty((n,), dtype=torch.uint8, device=_dev())
    dd = 0 if depth is None else L.dtype_code(depth)
    L.check(lib.gp_lift_mean_idmap(L.ptr(coords), L.dtype_code(coords), n, L.ptr(w2c), L.ptr(intr), L.ptr(depth), dd,
                                   V, H, W, cut, float(vis_thres), L.ptr(keep), L.ptr(idmap_u8), L.ptr(table),
                                   L.dtype_code(table), Q, D, L.ptr(out), out.stride(0), L.ptr(seen), L.stream()))
    L.count(1)
    return out, seen


def expand_idmap(idmap_u8, table_f32, dtype=torch.bfloat16, out=None):
    """feat[v,h,w,:] = table[v, idmap[v,h,w], :] as a dense channels-last map (the 2-D VLM stand-in)."""
    lib = L.load()
    V, H, W = idmap_u8.shape
    Q, D = table_f32.shape[1], table_f32.shape[2]
    if out is None:
        out = torch.empty((V, H, W, D), dtype=dtype, device=_dev())
    L.check(lib.gp_expand_idmap(L.ptr(idmap_u8), L.ptr(table_f32), V, H, W, Q, D, L.ptr(out), L.dtype_code(out),
                                L.stream()))
    L.count(1)
    return out


def lift_mean_pix(pix, keep, feat, H, W, out=None, out_stride=None):
    lib = L.load()
    V, n = pix.shape
    D = feat.shape[-1]
    if out is None:
        out_stride = out_stride or D
        out = _new((n, out_stride), torch.float32, "lifted")
    seen = torch.empty((n,), dtype=torch.uint8, device=_dev())
    L.check(lib.gp_lift_mean_pix(L.ptr(pix), n, V, H, W, L.ptr(keep), L.ptr(feat), L.dtype_code(feat), D, L.ptr(out),
                                 out.stride(0), L.ptr(seen), L.stream()))
    L.count(1)
    return out, seen


def lift_consensus(pix, keep, idmap_u8, table_f32, text, logit_scale, xyz_f32, cell=0.05, out=None, out_stride=None):
    """A5 lifting from the compact 2-D form (models/affinity_module.py:584-685).  Returns (out [n,stride] f32,
    seen [n] u8); never-seen rows are zero (fill them with nn1_fill + copy_rows)."""
    lib = L.load()
    V, n = pix.shape
    _, H, W = idmap_u8.shape
    Q, D = table_f32.shape[1], table_f32.shape[2]
    C = text.shape[0]
    if out is None:
        out_stride = out_stride or D
        out = _new((n, out_stride), torch.float32, "lifted")
    seen = torch.empty((n,), dtype=torch.uint8, device=_dev())
    nb = lib.gp_lift_consensus_workspace_bytes(n, V, Q, D, C)
    ws = _ws(nb, "lift_consensus")
    L.check(lib.gp_lift_consensus(L.ptr(pix), n, V, H, W, L.ptr(keep), L.ptr(idmap_u8), L.ptr(table_f32), Q, D,
                                  L.ptr(text), C, float(logit_scale), L.ptr(xyz_f32), float(cell), L.ptr(out),
                                  out.stride(0), L.ptr(seen), L.ptr(ws), nb, L.stream()))
    L.count(4 + 10 * V)
    return out, seen


def nn1_fill(xyz, seen, cell=0.05):
    """Nearest seen point of every unseen point: nn [N] int32 (nn[i]=i when seen)."""
    lib = L.load()
    n = xyz.shape[0]
    nn = _new((n,), torch.int32, "nn1")
    nb = lib.gp_nn1_fill_workspace_bytes(n)
    ws = _ws(nb, "nn1_fill")
    L.check(lib.gp_nn1_fill(L.ptr(xyz), n, L.ptr(seen), float(cell), L.ptr(nn), L.ptr(ws), nb, L.stream()))
    L.count(8)
    return nn


def copy_rows(rows, D, nn):
    lib = L.load()
    L.check(lib.gp_copy_rows(L.ptr(rows), rows.shape[0], D, rows.stride(0), L.ptr(nn), L.stream()))
    L.count(1)
    return rows


def voxelize(coords, xform):
    """F1 on the device.  coords [n,3] f32/f64 CUDA, xform [3,4] float64 (host array or CUDA tensor).
    Returns (vox [m,3] int32, inds [m] int32, inds_reconstruct [n] int32, extent tuple); one host sync for m."""
    lib = L.load()
    n = coords.shape[0]
    xf = torch.as_tensor(xform, dtype=torch.float64).reshape(12).to(_dev()).contiguous()
    vox = torch.empty((n, 3), dtype=torch.int32, device=_dev())
    inds = torch.empty((n,), dtype=torch.int32, device=_dev())
    rec = torch.empty((n,), dtype=torch.int32, device=_dev())
    meta = torch.zeros((4,), dtype=torch.int64, device=_dev())   # m | extent as int32[3] in the next 12 bytes
    nb = lib.gp_voxelize_workspace_bytes(n)
    ws = _ws(nb, "voxelize")
    ext_ptr = ctypes.c_void_p(meta.data_ptr() + 8)
    L.check(lib.gp_voxelize(L.ptr(coords), L.dtype_code(coords), n, L.ptr(xf), L.ptr(vox), L.ptr(inds), L.ptr(rec),
                            L.ptr(meta), ext_ptr, L.ptr(ws), nb, L.stream()))
    L.count(11)
    host = meta.cpu()
    m = int(host[0])
    ext = host[1:3].view(torch.int32)[:3].tolist()
    return vox[:m], inds[:m], rec, tuple(int(e) for e in ext)


class VoxelIndex:
    """Hilbert-cell-sorted voxel order + cell lists (device workspace)."""

    def __init__(self, vox_i32, extent=None):
        lib = L.load()
        self.m = vox_i32.shape[0]
        if extent is None:
            extent = (vox_i32.max(0).values + 1).tolist() if self.m else [1, 1, 1]
        self.ex, self.ey, self.ez = (int(e) for e in extent)
        nb = lib.gp_voxel_index_workspace_bytes(self.m, self.ex, self.ey, self.ez)
        if nb == 0:
            raise L.GeoPurifyError(f"voxel index: unsupported sizes m={self.m} extent={extent}")
        self.nbytes = nb
        self.ws = _new((max(int(nb), 256),), torch.uint8, "index_ws") if not _GUARD else _ws(nb)
        self.order = _new((self.m,), torch.int32, "index_order")
        self.rank = _new((self.m,), torch.int32, "index_rank")
        L.check(lib.gp_voxel_index_build(L.ptr(vox_i32, torch.int32), self.m, self.ex, self.ey, self.ez,
                                         L.ptr(self.order), L.ptr(self.rank), L.ptr(self.ws), nb, L.stream()))
        L.count(8)

    def check(self):
        """Synchronising validity check: raises if a voxel coordinate was outside [0, extent) (the build never writes
        out of bounds for such input, but kernel map and kNN would be meaningless)."""
        L.check(L.load().gp_voxel_index_check(L.ptr(self.ws), self.m, self.ex, self.ey, self.ez, L.stream()))
        return self

    def kernel_map(self):
        lib = L.load()
        out = _new((self.m, 27), torch.int32, "nbr27")
        L.check(lib.gp_kernel_map(L.ptr(self.ws), self.m, self.ex, self.ey, self.ez, L.ptr(out), L.stream()))
        L.count(1)
        return out

    def knn(self, k, want_orig=True, want_sorted=True):
        lib = L.load()
        no = _new((self.m, k), torch.int32, "knn_orig") if want_orig else None
        ns = _new((self.m, k), torch.int32, "knn_sorted") if want_sorted else None
        L.check(lib.gp_knn_grid(L.ptr(self.ws), self.m, self.ex, self.ey, self.ez, k, L.ptr(no), L.ptr(ns), L.stream()))
        L.count(1)
        return no, ns


def knn_points(xyz_f32, k, min_cell=1e-3):
    """F4: exact K nearest neighbours of raw float32 points, self excluded, ascending (float32 d2, index)
    (faiss search on scene_coords, models/affinity_module.py:1159-1165).  Returns [n,k] int32."""
    lib = L.load()
    n = xyz_f32.shape[0]
    out = torch.empty((n, k), dtype=torch.int32, device=_dev())
    nb = lib.gp_knn_points_workspace_bytes(n)
    ws = _ws(nb, "knn_points")
    L.check(lib.gp_knn_points(L.ptr(xyz_f32), n, k, float(min_cell), L.ptr(out), L.ptr(ws), nb, L.stream()))
    L.count(9)
    return out


def mine_contrastive(feat, anchors_i32, nbr_i32, n_macro=48, n_micro=15):
    """F4: positives and hybrid negatives of `sample_contrastive_pairs_hybrid` (models/affinity_module.py:1099-1136)
    for the given anchors.  Returns (positive [A] int32, negatives [A, n_macro + n_micro] int32)."""
    lib = L.load()
    n, d = feat.shape
    a, k = anchors_i32.shape[0], nbr_i32.shape[1]
    pos = torch.empty((a,), dtype=torch.int32, device=_dev())
    neg = torch.empty((a, n_macro + n_micro), dtype=torch.int32, device=_dev())
    nb = lib.gp_mine_workspace_bytes(n, d, a, n_macro)
    if nb == 0:
        raise L.GeoPurifyError(f"gp_mine_workspace_bytes: bad sizes n={n} D={d} A={a}")
    ws = _ws(nb, "mine")
    L.check(lib.gp_mine_contrastive(L.ptr(feat), feat.stride(0), d, n, L.ptr(anchors_i32), a, L.ptr(nbr_i32), k,
                                    n_macro, n_micro, L.ptr(pos), L.ptr(neg), L.ptr(ws), nb, L.stream()))
    L.count(6)
    return pos, neg


class ScatterPlan:
    """CSR of points per destination row, reused for every scatter_mean with the same index."""

    def __init__(self, p2v_i32, m):
        lib = L.load()
        self.n, self.m = p2v_i32.shape[0], int(m)
        nb = lib.gp_scatter_mean_workspace_bytes(self.n, self.m)
        self.ws = _new((max(int(nb), 256),), torch.uint8, "scatter_ws") if not _GUARD else _ws(nb)
        L.check(lib.gp_scatter_mean_prepare(L.ptr(p2v_i32, torch.int32), self.n, self.m, L.ptr(self.ws), nb, L.stream()))
        L.count(6)

    def mean(self, src, D, out, col0=0):
        lib = L.load()
        L.check(lib.gp_scatter_mean(L.ptr(src), src.stride(0), D, L.ptr(self.ws), self.n, self.m, L.ptr(out),
                                    out.stride(0), col0, L.stream()))
        L.count(1)
        return out


def affinity_softmax(embed, nbr_sorted, sharpen=20.0):
    lib = L.load()
    m, k = nbr_sorted.shape
    w = _new((m, k), torch.float32, "weights")
    L.check(lib.gp_affinity_softmax(L.ptr(embed, torch.float32), embed.shape[1], L.ptr(nbr_sorted, torch.int32), m, k,
                                    float(sharpen), L.ptr(w),
                                    L.stream()))
    L.count(1)
    return w


def pool(f0, D, nbr_sorted, weights, T, tensor_core=None):
    """T applications of the affinity matrix; returns a [m,D] tensor (or f0 itself when T == 0).
    D in (256, 512) runs on the tensor cores (per-scene plan + tcgen05 block GEMM, one 256-column pass per work item);
    other widths one warp per row."""
    lib = L.load()
    m, k = nbr_sorted.shape
    if T == 0:
        return f0
    if tensor_core is None:
        # the per-scene plan costs about as much as one row-wise application: it pays off from T = 2
        # (profiles/r01_sweep_pool_knn.jsonl)
        tensor_core = D in (256, 512) and T >= 2
    if tensor_core:
        nbp = lib.gp_pool_plan_bytes(m, k)
        plan = _ws(nbp, "pool_plan")
        L.check(lib.gp_pool_plan(L.ptr(nbr_sorted, torch.int32), L.ptr(weights, torch.float32), m, k, L.ptr(plan), nbp,
                                 L.stream()))
        nbw = lib.gp_pool_tc_workspace_bytes(m)
        ws = _ws(nbw, "pool_ws")
        out = torch.empty((m, D), dtype=torch.float32, device=_dev())
        L.check(lib.gp_pool_tc(L.ptr(f0), f0.stride(0), D, L.ptr(plan), L.ptr(nbr_sorted), L.ptr(weights), m, k, T,
                               L.ptr(out), L.ptr(ws), nbw, L.stream()))
        L.count(2 + 2 * T)
        return out
    a = torch.empty((m, D), dtype=torch.float32, device=_dev())
    b = torch.empty((m, D), dtype=torch.float32, device=_dev()) if T > 1 else a
    outp = ctypes.c_void_p()
    L.check(lib.gp_pool(L.ptr(f0, torch.float32), f0.stride(0), D, L.ptr(nbr_sorted, torch.int32),
                        L.ptr(weights, torch.float32), m, k, T, L.ptr(a), L.ptr(b),
                        ctypes.byref(outp), L.stream()))
    L.count(T)
    return a if outp.value == a.data_ptr() else b


def gather_rows(rows, D, index_i32):
    lib = L.load()
    n = index_i32.shape[0]
    out = torch.empty((n, D), dtype=torch.float32, device=_dev())
    L.check(lib.gp_gather_rows(L.ptr(rows, torch.float32), rows.stride(0), D, L.ptr(index_i32, torch.int32), n,
                               L.ptr(out), L.stream()))
    L.count(1)
    return out


def logits_argmax(feat, D, text, logit_scale, want_logits=False, want_zero=True, tensor_core=None):
    """A12: (label [rows] int32, zero [rows] uint8 | None, logits [rows,C] | None).  Without `want_logits` the labels
    come from the tcgen05 GEMM with the fused arg-max epilogue (identical labels, see gp_logits_argmax_tc)."""
    lib = L.load()
    rows, C = feat.shape[0], text.shape[0]
    label = torch.empty((rows,), dtype=torch.int32, device=_dev())
    zero = torch.empty((rows,), dtype=torch.uint8, device=_dev()) if want_zero else None
    if tensor_core is None:
        tensor_core = (not want_logits) and C <= 256 and D % 128 == 0 and float(logit_scale) > 0 and rows > 0
    if tensor_core:
        nb = lib.gp_logits_tc_workspace_bytes(rows, C, D)
        ws = _ws(nb, "logits_tc")
        L.check(lib.gp_logits_argmax_tc(L.ptr(feat), feat.stride(0), D, rows, L.ptr(text), C, float(logit_scale),
                                        L.ptr(label), L.ptr(zero), L.ptr(ws), nb, L.stream()))
        L.count(6)
        return label, zero, None
    logits = torch.empty((rows, C), dtype=torch.float32, device=_dev()) if want_logits else None
    nb = lib.gp_logits_workspace_bytes(C, D)
    ws = _ws(nb, "logits")
    L.check(lib.gp_logits_argmax(L.ptr(feat), feat.stride(0), D, rows, L.ptr(text), C, float(logit_scale), L.ptr(label),
                                 L.ptr(zero), L.ptr(logits), L.ptr(ws), nb, L.stream()))
    L.count(2)
    return label, zero, logits


def iou_hist(label_rows, index, target, ignore, C, hist=None, want_pred=False):
    lib = L.load()
    n = target.shape[0]
    if hist is None:
        hist = torch.zeros((3, C), dtype=torch.int64, device=_dev())
    if ignore is None:       # no ignore ids
        ignore = torch.empty((0,), dtype=torch.int64, device=_dev())
    pred = torch.empty((n,), dtype=torch.int64, device=_dev()) if want_pred else None
    L.check(lib.gp_iou_hist(L.ptr(label_rows, torch.int32), L.ptr(index, torch.int32), L.ptr(target, torch.int64), n,
                            L.ptr(ignore, torch.int64), ignore.shape[0], C, L.ptr(pred), L.ptr(hist, torch.int64),
                            L.stream()))
    L.count(1)
    return hist, pred
