"""CPU oracle for GeoPurify's inference hot path -- TEST INFRASTRUCTURE, NOT PRODUCT.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import this module.  The product path (`geopurify_b200/`) never does and fails
loudly when the CUDA library is missing.

Each function restates one row of SURVEY.md §8(a) in numpy / torch-CPU, citing the
reference lines it follows (paths relative to the reference tree).

Pinning status
  * A1/A2 (projection, occlusion, z-buffer depth): PINNED -- checked bit-for-bit against
    the reference's own `PointCloudToImageMapper` / `PointCloudToImageMappermatterport`
    (imported from the reference tree by `tests/golden/make_golden.py`; vectors committed
    under `tests/golden/`).
  * F1 (voxeliser) and the A12 metric: PINNED by the reference's own `Voxelizer.voxelize` and
    `intersectionAndUnionGPU` (`tests/golden/make_golden_voxel.py`, `make_golden_metric.py`).
  * A4-A12, F3: PINNED against runs of the reference's own `SonataXAffinityTrainer.evaluate_scene`,
    `lift_lseg_features`, `lift_xdecoder_features` and `run/validation.py:validate`
    (`tests/golden/make_golden_scene.py`; vectors `scene_lseg.npz`, `scene_xdecoder.npz`,
    `validate_two_scenes.npz`).  In those runs MinkowskiEngine, faiss.IndexFlatL2 and
    torch_scatter.scatter_mean -- neither vendored nor installable here -- are stood in by
    definition stubs (`tests/golden/ref_stubs/`): scatter_mean = sum / count; ME conv = sum
    over occupied 3^3 offsets of W_d^T x, offsets x fastest, no bias; IndexFlatL2 = exact L2,
    ties by lowest index as the north star states.  The arithmetic INSIDE those three
    operators therefore rests on their documentation; everything around them is the
    reference's code.  sklearn `KDTree` is the real library.  The offset convention of the
    stubbed convolution is pinned by `test_sparse_conv_offset_convention_delta_response`.
  * C helpers (oracle/libgp_oracle.so, built by oracle/Makefile): `knn_ref.c` (OpenMP brute-force
    kNN, checked against the numpy restatement) and `spmm_ref.c` (OpenMP CSR SpMM, checked BIT FOR
    BIT against `torch.sparse.mm`, the reference's own pooling call).  They exist so that the
    full-size parity test (N = 150 000, K = 96, T = 19) and the CPU arm of bench.py run the real
    workload in seconds; `pool_impl="torch"` keeps the literal `torch.sparse.mm` path.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

# ----------------------------------------------------------------------------- A1 / A2


def mapper_intrinsics(intrinsics, image_dim):
    """`PointCloudToImageMapper.__init__` rescale (models/utils/fusion_util.py:90-96)."""
    k = np.array(intrinsics).copy()
    sx = image_dim[0] / (k[0, 2] * 2)
    sy = image_dim[1] / (k[1, 2] * 2)
    k[0, 0] *= sx
    k[1, 1] *= sy
    k[0, 2] = image_dim[0] / 2
    k[1, 2] = image_dim[1] / 2
    return k


def project_view(world_to_camera_T, coords, depth, intrinsic, image_dim, vis_thres, cut, splat=False):
    """`PointCloudToImageMapper.compute_mapping` (models/utils/fusion_util.py:99-147).

    `world_to_camera_T` is W2C^T (models/scene/camera.py:87); `intrinsic` is the already
    rescaled matrix.  Returns (mapping [N,3] int64 = (row, col, mask), depth_used).
    `splat=True` is the string-depth branch (:126-130) restated with an order-independent min.
    """
    n = coords.shape[0]
    mapping = np.zeros((3, n), dtype=int)
    coords_new = np.concatenate([coords, np.ones([n, 1])], axis=1).T          # :111
    p = np.matmul(world_to_camera_T.T, coords_new)                            # :114
    with np.errstate(all="ignore"):
        p[0] = (p[0] * intrinsic[0][0]) / p[2] + intrinsic[0][2]              # :115
        p[1] = (p[1] * intrinsic[1][1]) / p[2] + intrinsic[1][2]              # :116
        pi = np.round(p).astype(int)                                          # :117
    inside = ((pi[0] >= cut) * (pi[1] >= cut) * (pi[0] < image_dim[0] - cut) * (pi[1] < image_dim[1] - cut))
    if splat:                                                                 # :126-130
        d = np.full(image_dim[1] * image_dim[0], 999999.0)
        sel = inside & (p[2] > 0.2)
        np.minimum.at(d, pi[1][sel] * image_dim[0] + pi[0][sel], p[2][sel])
        depth = d.reshape(image_dim[1], image_dim[0])
    if depth is not None:                                                     # :132-138
        depth_cur = depth[pi[1][inside], pi[0][inside]]
        occl = np.abs(depth[pi[1][inside], pi[0][inside]] - p[2][inside]) <= vis_thres * depth_cur
        inside[inside == True] = occl  # noqa: E712
    else:                                                                     # :139-141
        inside = (p[2] > 0) * inside
    mapping[0][inside] = pi[1][inside]
    mapping[1][inside] = pi[0][inside]
    mapping[2][inside] = 1
    return mapping.T, depth


def project_view_matterport(camera_to_world, coords, depth, intrinsic, image_dim, vis_thres, cut):
    """`PointCloudToImageMappermatterport.compute_mapping` (models/utils/fusion_util.py:45-82)."""
    n = coords.shape[0]
    mapping = np.zeros((3, n), dtype=int)
    coords_new = np.concatenate([coords, np.ones([n, 1])], axis=1).T
    world_to_camera = np.linalg.inv(camera_to_world)                          # :60
    p = np.matmul(world_to_camera, coords_new)
    with np.errstate(all="ignore"):
        p[0] = (p[0] * intrinsic[0][0]) / p[2] + intrinsic[0][2]
        p[1] = (p[1] * intrinsic[1][1]) / p[2] + intrinsic[1][2]
        pi = np.round(p).astype(int)
    inside = ((pi[0] >= cut) * (pi[1] >= cut) * (pi[0] < image_dim[0] - cut) * (pi[1] < image_dim[1] - cut))
    if depth is not None:
        depth_cur = depth[pi[1][inside], pi[0][inside]]
        occl = np.abs(depth[pi[1][inside], pi[0][inside]] - p[2][inside]) <= vis_thres * depth_cur
        inside[inside == True] = occl  # noqa: E712
    else:
        inside = (p[2] > 0) * inside
    mapping[0][inside] = pi[1][inside]
    mapping[1][inside] = pi[0][inside]
    mapping[2][inside] = 1
    return mapping.T


# ----------------------------------------------------------------------------- A3
def accept_views(mappings, min_visible=400, max_visible=None):
    """View acceptance + renumbering (dataset/data_loader_ablation.py:254-288; collate :430-459).

    `mappings` is [V,N,3].  Returns kept view ids, and per kept view (rows, cols, point ids)
    in point order -- the `x_label` / `y_label` / `mask_2d` triple of the reference.
    """
    kept, rows, cols, pids = [], [], [], []
    for v in range(mappings.shape[0]):
        m = mappings[v][:, 2]
        s = int(m.sum())
        if s == 0 or s < min_visible or (max_visible is not None and s > max_visible):
            continue
        idx = np.nonzero(m == 1)[0]
        kept.append(v)
        rows.append(mappings[v][idx, 0])
        cols.append(mappings[v][idx, 1])
        pids.append(idx)
    return kept, rows, cols, pids


# ----------------------------------------------------------------------------- A4
def nn1_fill_index(seen_xyz, query_xyz):
    """k=1 KDTree query as used for every hole fill (models/affinity_module.py:445-449, 619-625, 693-696)."""
    from sklearn.neighbors import KDTree
    _, idx = KDTree(seen_xyz).query(query_xyz, k=1)
    return idx.flatten()


def lift_mean(n_points, scene_xyz, view_features, rows, cols, pids, feat_dim=512):
    """LSeg-style lifting: gather + mean over views, 1-NN fill (models/affinity_module.py:404-449).

    `view_features(i)` returns the channels-last [H,W,D] float32 map of kept view i; the
    reference indexes a [D,H,W] map as feat[:, x, y] (:426), which reads the same element.
    Sum is taken view by view in float32 (`index_add_`, :428).
    """
    s = torch.zeros((n_points, feat_dim), dtype=torch.float32)
    cnt = torch.zeros((n_points, 1), dtype=torch.float32)
    for i in range(len(pids)):
        vf = view_features(i)
        if isinstance(vf, tuple):      # (idmap [H,W], table [Q,D]): same element, no dense map built
            idm, tab = vf
            lifted = torch.as_tensor(tab).float()[torch.as_tensor(idm[rows[i], cols[i]].astype(np.int64))]
        else:
            lifted = torch.as_tensor(vf).float()[torch.as_tensor(rows[i]), torch.as_tensor(cols[i])]
        pi = torch.as_tensor(pids[i])
        s.index_add_(0, pi, lifted)
        cnt.index_add_(0, pi, torch.ones((len(pi), 1), dtype=torch.float32))
    cnt[cnt == 0] = 1e-6                                                       # :435
    out = s / cnt                                                              # :436
    seen = cnt.squeeze(1) > 1e-5                                               # :438
    if (~seen).any() and seen.any():                                           # :441-449
        xyz = np.asarray(scene_xyz)
        idx = nn1_fill_index(xyz[seen.numpy()], xyz[(~seen).numpy()])
        out[~seen] = out[seen][torch.as_tensor(idx)]
    return out, seen


# ----------------------------------------------------------------------------- A5
def lift_consensus(n_points, scene_xyz, view_point_feats, pids, text, logit_scale, feat_dim=512):
    """X-Decoder-style lifting (models/affinity_module.py:584-696).

    `view_point_feats[i]` is the [n_i, D] per-point feature of kept view i BEFORE the per-view
    hole fill: the covering mask's embedding, or zeros (:584-592).
    """
    text_n = F.normalize(torch.as_tensor(text).float(), dim=-1)               # :629
    xyz = torch.as_tensor(np.asarray(scene_xyz)).float()
    feats, logits = [], []
    for i in range(len(pids)):
        f = torch.as_tensor(view_point_feats[i]).float().clone()
        rs = torch.sum(f, dim=1)
        zero = rs == 0                                                         # :606, :614-617
        if zero.any() and (~zero).any():
            p = xyz[torch.as_tensor(pids[i])]
            idx = nn1_fill_index(p[~zero].numpy(), p[zero].numpy())
            true_idx = torch.where(~zero)[0]
            f[zero] = f[true_idx[torch.as_tensor(idx)]]                        # :623-625
        f = F.normalize(f, dim=-1)                                             # :628
        feats.append(f)
        logits.append(logit_scale * (f @ text_n.t()))                          # :630
    # regroup per point, views in ascending order (:633-638)
    nv = torch.zeros(n_points, dtype=torch.long)
    for pi in pids:
        nv[torch.as_tensor(pi)] += 1
    mmax = int(nv.max()) if n_points else 0
    c = text_n.shape[0]
    pf = torch.zeros(n_points, mmax, feat_dim)
    pl = torch.zeros(n_points, mmax, c)
    valid = torch.zeros(n_points, mmax, dtype=torch.bool)
    slot = torch.zeros(n_points, dtype=torch.long)
    for i, pi in enumerate(pids):
        pi = torch.as_tensor(pi)
        pf[pi, slot[pi]] = feats[i]
        pl[pi, slot[pi]] = logits[i]
        valid[pi, slot[pi]] = True
        slot[pi] += 1
    out = torch.zeros(n_points, feat_dim)
    seen = nv > 0
    if seen.any():
        k = min(mmax, 3)                                                       # :672 (per chunk; same value when the chunk holds a >=3-view point)
        avg = pl.sum(1) / valid.sum(1, keepdim=True).clamp(min=1)              # :673-675
        cidx = torch.argmax(avg, dim=1)                                        # :676
        score = torch.gather(pl, 2, cidx.view(-1, 1, 1).expand(-1, mmax, -1)).squeeze(-1)
        score = score.masked_fill(~valid, -torch.inf)                          # :678
        top, ti = torch.topk(score, k=k, dim=1)                                # :679
        tf = torch.gather(pf, 1, ti.unsqueeze(-1).expand(-1, -1, feat_dim))
        w = F.softmax(top, dim=1)                                              # :682
        fused = (tf * w.unsqueeze(-1)).sum(1)                                  # :683
        out[seen] = fused[seen]
    if seen.any() and (~seen).any():                                           # :687-696
        idx = nn1_fill_index(xyz[seen].numpy(), xyz[~seen].numpy())
        out[~seen] = out[seen][torch.as_tensor(idx)]
    return out, seen


# ----------------------------------------------------------------------------- F3
def mask_partition(mask_logits, cls_logits, score_thresh=0.0):
    """Panoptic-style mask selection of one view (models/affinity_module.py:544-570), restated with its Python loop.
    mask_logits [Q,H,W], cls_logits [Q,C+1].  Returns (final_keep_indices in ORIGINAL query numbering,
    final_masks list of bool [H,W], idmap [H,W] int64 with Q where no final mask covers the pixel)."""
    mask_logits = torch.as_tensor(mask_logits).float()
    cls_logits = torch.as_tensor(cls_logits).float()
    q_total = mask_logits.shape[0]
    scores, labels = F.softmax(cls_logits, dim=-1)[..., :-1].max(-1)            # :544
    keep = scores > score_thresh                                                # :546
    cur_scores = scores[keep]
    cur_masks = mask_logits[keep].sigmoid()                                     # :549
    kept_ids = torch.nonzero(keep).flatten()
    idmap = torch.full(mask_logits.shape[1:], q_total, dtype=torch.long)
    final_keep, final_masks = [], []
    if cur_masks.shape[0] == 0:
        return final_keep, final_masks, idmap
    cur_prob_masks = cur_scores.view(-1, 1, 1) * cur_masks                      # :557
    cur_mask_ids = cur_prob_masks.argmax(0)                                     # :558
    for k in range(cur_scores.shape[0]):                                        # :562-570
        mask_area = (cur_mask_ids == k).sum().item()
        original_area = (cur_masks[k] >= 0.5).sum().item()
        mask = (cur_mask_ids == k) & (cur_masks[k] >= 0.5)
        if mask_area > 0 and original_area > 0 and mask.sum().item() > 0:
            final_keep.append(int(kept_ids[k]))
            final_masks.append(mask)
            idmap[mask] = int(kept_ids[k])
    return final_keep, final_masks, idmap


# ----------------------------------------------------------------------------- F1
def fnv_hash_vec(arr):
    """dataset/voxelization_utils.py:6-18: uint64 FNV-1-style key, multiply by the prime then xor, per column."""
    arr = np.asarray(arr).astype(np.uint64)
    h = np.full(arr.shape[0], np.uint64(14695981039346656037), dtype=np.uint64)
    with np.errstate(over="ignore"):
        for j in range(arr.shape[1]):
            h = h * np.uint64(1099511628211)
            h = np.bitwise_xor(h, arr[:, j])
    return h


def voxelize(coords, voxel_size=0.02, rigid_transformation=None):
    """Voxelizer.voxelize without clipping / random augmentation (dataset/voxelizer.py:104-121) followed by
    sparse_quantize(return_index=True) (dataset/voxelization_utils.py:78-96).
    Returns (coords_aug[inds] [M,3] float64, inds [M], inds_reconstruct [N])."""
    coords = np.asarray(coords)
    if rigid_transformation is None:
        rigid_transformation = np.eye(4)
        np.fill_diagonal(rigid_transformation[:3, :3], 1 / voxel_size)                 # voxelizer.py:61-63
    homo = np.hstack((coords, np.ones((coords.shape[0], 1), dtype=coords.dtype)))      # :107-109
    aug = np.floor(homo @ rigid_transformation.T[:, :3])                               # :110
    aug = np.floor(aug - aug.min(0))                                                   # :112-116
    key = fnv_hash_vec(np.floor(aug / np.array([1, 1, 1])))                            # voxelization_utils.py:76-81
    _, inds, inv = np.unique(key, return_index=True, return_inverse=True)              # :93
    return aug[inds], inds, inv.astype(np.int64)


# ----------------------------------------------------------------------------- A6
def scatter_mean(src, index, m=None):
    """torch_scatter.scatter_mean(src, index, dim=0) (models/affinity_module.py:1524-1535):
    float32 sum per voxel divided by the count (clamped to 1)."""
    src = torch.as_tensor(src).float()
    index = torch.as_tensor(index).long()
    m = int(index.max()) + 1 if m is None else m
    out = torch.zeros((m, src.shape[1]), dtype=torch.float32)
    out.index_add_(0, index, src)
    cnt = torch.bincount(index, minlength=m).clamp(min=1).float()
    return out / cnt[:, None]


# ----------------------------------------------------------------------------- A7 / A8
def kernel_map(voxel_coords):
    """3^3 neighbour table: nbr[i, o] = index of voxel at coords[i] + delta_o or -1, with
    o = (dx+1) + 3(dy+1) + 9(dz+1) (MinkowskiEngine region order, x fastest)."""
    c = np.asarray(voxel_coords).astype(np.int64)
    m = c.shape[0]
    lut = {(int(x), int(y), int(z)): i for i, (x, y, z) in enumerate(c)}
    nbr = np.full((m, 27), -1, dtype=np.int64)
    for o in range(27):
        dx, dy, dz = o % 3 - 1, (o // 3) % 3 - 1, o // 9 - 1
        for i in range(m):
            nbr[i, o] = lut.get((c[i, 0] + dx, c[i, 1] + dy, c[i, 2] + dz), -1)
    return nbr


def kernel_map_fast(voxel_coords):
    """Vectorised `kernel_map` (sorted-key lookup) for larger cases."""
    c = np.asarray(voxel_coords).astype(np.int64) + 1
    ext = c.max(0) + 2
    key = (c[:, 2] * ext[1] + c[:, 1]) * ext[0] + c[:, 0]
    order = np.argsort(key, kind="stable")
    skey = key[order]
    m = c.shape[0]
    nbr = np.full((m, 27), -1, dtype=np.int64)
    for o in range(27):
        dx, dy, dz = o % 3 - 1, (o // 3) % 3 - 1, o // 9 - 1
        q = key + (dz * ext[1] + dy) * ext[0] + dx
        pos = np.searchsorted(skey, q)
        pos = np.minimum(pos, m - 1)
        hit = skey[pos] == q
        nbr[hit, o] = order[pos[hit]]
    return nbr


def sparse_conv3(x, w, nbr):
    """ME.MinkowskiConvolution(kernel_size=3, stride=1, bias=False) (models/affinity_module.py:36-38, 56):
    out[i] = sum_o W_o^T x[nbr[i,o]] over occupied offsets, o ascending, float32."""
    out = torch.zeros((x.shape[0], w.shape[2]), dtype=torch.float32)
    nbr_t = torch.as_tensor(nbr)
    for o in range(27):
        rows = torch.where(nbr_t[:, o] >= 0)[0]
        if rows.numel():
            out.index_add_(0, rows, x[nbr_t[rows, o]] @ w[o])
    return out


def batch_norm(x, sd, name, eps=1e-5):
    """ME.MinkowskiBatchNorm in eval mode == BatchNorm1d on the feature matrix."""
    return F.batch_norm(x, sd[name + ".bn.running_mean"], sd[name + ".bn.running_var"],
                        sd[name + ".bn.weight"], sd[name + ".bn.bias"], False, 0.0, eps)


def student_forward(x, voxel_coords, state, nbr=None):
    """`AffinityPredictor.forward` + F.normalize (models/affinity_module.py:33-72, 1546-1547)."""
    sd = {k: torch.as_tensor(v).float() for k, v in state.items()}
    if nbr is None:
        nbr = kernel_map_fast(voxel_coords)
    x = torch.as_tensor(x).float()
    out = F.relu(batch_norm(sparse_conv3(x, sd["input_layer.0.kernel"], nbr), sd, "input_layer.1"))
    for b in range(4):
        idn = out
        y = F.relu(batch_norm(sparse_conv3(out, sd[f"res_blocks.{b}.conv1.kernel"], nbr), sd, f"res_blocks.{b}.norm1"))
        y = batch_norm(sparse_conv3(y, sd[f"res_blocks.{b}.conv2.kernel"], nbr), sd, f"res_blocks.{b}.norm2")
        out = F.relu(y + idn)
    emb = out @ sd["output_layer.kernel"]
    return F.normalize(emb, p=2, dim=1)


# ----------------------------------------------------------------------------- A9
def knn(voxel_coords, k, chunk=2048):
    """faiss.IndexFlatL2.search(x, k+1)[:, 1:] (models/affinity_module.py:1551-1557) with the
    north star's tie rule: ascending (d^2, index); self (d^2 = 0, unique coords) dropped."""
    c = np.asarray(voxel_coords).astype(np.int64)
    m = c.shape[0]
    assert m >= k + 1, "reference needs M >= K+1 (faiss pads with -1 otherwise)"
    out = np.empty((m, k), dtype=np.int64)
    idx = np.arange(m, dtype=np.int64)
    for s in range(0, m, chunk):
        q = c[s:s + chunk]
        d2 = ((q[:, None, :] - c[None, :, :]) ** 2).sum(-1)
        key = d2 * m + idx[None, :]
        part = np.argpartition(key, k, axis=1)[:, :k + 1]
        pk = np.take_along_axis(key, part, 1)
        order = np.argsort(pk, axis=1)
        out[s:s + chunk] = np.take_along_axis(part, order, 1)[:, 1:]
    return out


def knn_c(voxel_coords, k):
    """Same contract as `knn`, through oracle/libgp_oracle.so (knn_ref.c, OpenMP brute force)."""
    import ctypes
    import os
    so = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libgp_oracle.so")
    if not os.path.exists(so):
        return knn(voxel_coords, k)
    lib = ctypes.CDLL(so)
    c = np.ascontiguousarray(np.asarray(voxel_coords).astype(np.int32))
    out = np.empty((c.shape[0], k), dtype=np.int64)
    rc = lib.gp_oracle_knn(c.ctypes.data_as(ctypes.c_void_p), ctypes.c_int64(c.shape[0]), ctypes.c_int(k),
                           out.ctypes.data_as(ctypes.c_void_p))
    assert rc == 0, "reference needs M >= K+1"
    return out


# ----------------------------------------------------------------------------- A10 / A11
def knn_points(xyz, k, chunk=512):
    """Training-side kNN on raw points (models/affinity_module.py:1159-1165): faiss IndexFlatL2 over float32
    coordinates, K+1 results with the self column dropped.  Restated as the direct float32 evaluation
    d2 = (dx*dx + dy*dy) + dz*dz (each numpy float32 operation rounds once), ascending (d2, index), self excluded."""
    x = np.ascontiguousarray(xyz, dtype=np.float32)
    n = x.shape[0]
    out = np.empty((n, k), dtype=np.int64)
    for s in range(0, n, chunk):
        q = x[s:s + chunk]
        dx = x[None, :, 0] - q[:, None, 0]
        dy = x[None, :, 1] - q[:, None, 1]
        dz = x[None, :, 2] - q[:, None, 2]
        d2 = (dx * dx + dy * dy) + dz * dz
        d2[np.arange(q.shape[0]), np.arange(s, s + q.shape[0])] = np.inf
        out[s:s + chunk] = np.argsort(d2, axis=1, kind="stable")[:, :k]
    return out


def mine_contrastive(feat, anchor_indices, neighbor_indices, n_neg=63, n_macro=48):
    """`sample_contrastive_pairs_hybrid` after its randperm (models/affinity_module.py:1112-1136), same torch calls.
    Returns (positive [A], negatives [A, n_neg], sim [A, N] with the anchor / positive entries at +inf)."""
    f = torch.as_tensor(feat).float()
    anchor_indices = torch.as_tensor(anchor_indices).long()
    neighbor_indices = torch.as_tensor(neighbor_indices).long()
    num_points = f.shape[0]
    sim = torch.einsum('ad,pd->ap', F.normalize(f[anchor_indices], p=2, dim=1), F.normalize(f, p=2, dim=1))   # :1113
    sim_pos = sim.clone()
    sim_pos.scatter_(1, anchor_indices.unsqueeze(1), float('-inf'))                                          # :1117
    positive = torch.argmax(sim_pos, dim=1)                                                                   # :1118
    n_micro = n_neg - n_macro
    ar = torch.arange(num_points)
    exclude = (ar.unsqueeze(0) == anchor_indices.unsqueeze(1)) | (ar.unsqueeze(0) == positive.unsqueeze(1))   # :1124
    sim[exclude] = float('inf')                                                                               # :1126
    _, macro = torch.topk(sim, k=n_macro, largest=False, dim=1)                                               # :1127
    nb = neighbor_indices[anchor_indices]                                                                     # :1129
    local = torch.gather(sim, 1, nb)                                                                          # :1130
    _, hardest = torch.topk(local, k=n_micro, largest=False, dim=1)                                           # :1131
    micro = torch.gather(nb, 1, hardest)                                                                      # :1132
    return positive, torch.cat([macro, micro], dim=1), sim


def affinity_weights(embed, nbr_idx, sharpen=20.0, chunk=16384):
    """cosine affinities + softmax(20 x) (models/affinity_module.py:1559-1572).  Rows are independent, so the
    [M*K,128] temporaries of the reference (6 GB each at M = 130k) are built `chunk` rows at a time."""
    e = torch.as_tensor(embed).float()
    idx = torch.as_tensor(nbr_idx).long()
    m, k = idx.shape
    out = torch.empty((m, k), dtype=torch.float32)
    for s in range(0, m, chunk):
        ii = idx[s:s + chunk]
        center = e[s:s + chunk].repeat_interleave(k, dim=0)
        neigh = e[ii.flatten()]
        a = torch.einsum("bd,bd->b", center, neigh)
        out[s:s + chunk] = F.softmax(a.view(ii.shape[0], k) * sharpen, dim=1)
    return out


def pool(feat, nbr_idx, weights, n_iter=19):
    """F <- A F, `n_iter` times, A = COO(rows, nbr, weights) (models/affinity_module.py:1575-1587)."""
    f = torch.as_tensor(feat).float()
    idx = torch.as_tensor(nbr_idx).long()
    m, k = idx.shape
    rows = torch.arange(m).repeat_interleave(k)
    a = torch.sparse_coo_tensor(torch.stack([rows, idx.flatten()]), torch.as_tensor(weights).flatten(), size=(m, m),
                                check_invariants=False)
    for _ in range(n_iter):
        f = torch.sparse.mm(a, f)
    return f


def _oracle_lib():
    import ctypes
    import os
    so = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libgp_oracle.so")
    return ctypes.CDLL(so) if os.path.exists(so) else None


def morton_order(voxel_coords):
    """Visiting order for `pool_c`: voxels sorted by the Morton code of their integer coordinates."""
    c = np.asarray(voxel_coords).astype(np.int64)
    c = c - c.min(0)
    key = np.zeros(c.shape[0], dtype=np.int64)
    for b in range(20):
        for a in range(3):
            key |= ((c[:, a] >> b) & 1) << (3 * b + a)
    return np.ascontiguousarray(np.argsort(key, kind="stable").astype(np.int64))


def pool_c(feat, nbr_idx, weights, n_iter=19, visit_order=None):
    """Same contract as `pool`, through oracle/libgp_oracle.so (spmm_ref.c: OpenMP CSR SpMM that accumulates every
    output row entry by entry in list order with fused multiply-adds, bit for bit what torch.sparse.mm returns).
    `visit_order` only changes the order in which the independent rows are computed (cache locality), never the
    result.  ~40x faster than torch.sparse.mm on the host at M = 130k: the full-size parity test and the CPU arm of
    bench.py run the real 19 applications in a few seconds."""
    import ctypes
    lib = _oracle_lib()
    if lib is None or not hasattr(lib, "gp_oracle_spmm"):
        return pool(feat, nbr_idx, weights, n_iter)
    f = np.array(torch.as_tensor(feat).float().numpy(), dtype=np.float32, order="C", copy=True)   # ping-pong buffers: never the caller's
    idx = np.ascontiguousarray(np.asarray(nbr_idx, dtype=np.int64))
    w = np.ascontiguousarray(torch.as_tensor(weights).float().numpy())
    m, k = idx.shape
    g = np.empty_like(f)
    vo = None if visit_order is None else np.ascontiguousarray(np.asarray(visit_order, dtype=np.int64))
    assert vo is None or (vo.shape == (m,) and np.array_equal(np.sort(vo), np.arange(m))), "visit_order: a permutation"
    for _ in range(n_iter):
        rc = lib.gp_oracle_spmm(idx.ctypes.data_as(ctypes.c_void_p), w.ctypes.data_as(ctypes.c_void_p),
                                ctypes.c_int64(m), ctypes.c_int(k), f.ctypes.data_as(ctypes.c_void_p),
                                ctypes.c_int64(f.shape[1]), g.ctypes.data_as(ctypes.c_void_p),
                                None if vo is None else vo.ctypes.data_as(ctypes.c_void_p))
        assert rc == 0, "neighbour index out of range"
        f, g = g, f
    return torch.from_numpy(f)


# ----------------------------------------------------------------------------- A12
def classify(scene_features, text, logit_scale, yz=None):
    """run/validation.py:413-432.  `yz` = scene_coords[:, 1:4] (the (y,z) quirk) for the
    zero-row label fill; labels by `torch.max(.,1)[1]` (lowest index on ties)."""
    f = F.normalize(torch.as_tensor(scene_features).float(), dim=-1)
    t = F.normalize(torch.as_tensor(text).float(), dim=-1)
    logits = logit_scale * (f @ t.t())
    pred = torch.max(logits, 1)[1]
    unseen = torch.sum(f.abs(), dim=1) == 0
    if unseen.any() and (~unseen).any() and yz is not None:
        yz = np.asarray(yz)
        idx = nn1_fill_index(yz[(~unseen).numpy()], yz[unseen.numpy()])
        pred[unseen] = pred[torch.where(~unseen)[0][torch.as_tensor(idx)]]
    return pred, logits


def intersection_and_union(output, target, k, ignore_indexs=(255,)):
    """`intersectionAndUnionGPU` (util/util.py:160-177) on CPU tensors; float32 histograms."""
    output = torch.as_tensor(output).clone().view(-1)
    target = torch.as_tensor(target).view(-1)
    for ig in ignore_indexs:
        output[target == ig] = ig
    inter = output[output == target]
    ai = torch.histc(inter.float(), bins=k, min=0, max=k - 1)
    ao = torch.histc(output.float(), bins=k, min=0, max=k - 1)
    at = torch.histc(target.float(), bins=k, min=0, max=k - 1)
    return ai, ao + at - ai, at


def miou(inter_sum, union_sum, target_sum):
    """run/validation.py:513-523 on float32 numpy sums."""
    i = np.asarray(inter_sum, dtype=np.float32)
    u = np.asarray(union_sum, dtype=np.float32)
    t = np.asarray(target_sum, dtype=np.float32)
    iou = i / (u + 1e-10)
    acc = i / (t + 1e-10)
    return float(np.mean(iou)), float(np.mean(acc)), float(sum(i) / (sum(t) + 1e-10))


# ----------------------------------------------------------------------------- whole scene
def evaluate_scene(scene, state, k=96, n_iter=19, sharpen=20.0, min_visible=400, stages=None, pool_impl="auto"):
    """Stages 1a-4 on one synthetic `Scene` (LSeg-style mean lifting); returns a dict of every
    intermediate the parity tests compare.  `stages` (optional dict) receives wall seconds.
    `pool_impl`: "torch" = torch.sparse.mm as the reference calls it, "c" = spmm_ref.c (same accumulation order),
    "auto" = torch below 2M non-zeros, C above (torch.sparse.mm needs 7 s per application at M = 130k)."""
    import time
    t0 = time.perf_counter()

    def lap(name):
        nonlocal t0
        if stages is not None:
            t1 = time.perf_counter()
            stages[name] = stages.get(name, 0.0) + (t1 - t0)
            t0 = t1

    W, H = scene.image_dim
    kmap = mapper_intrinsics(scene.intrinsics, scene.image_dim)
    c64 = scene.coords.astype(np.float64)
    maps = np.stack([project_view(scene.w2c[v].T.copy(), c64, scene.depth[v], kmap, (W, H),
                                  scene.vis_thres, scene.cut)[0] for v in range(scene.n_views)])
    lap("project")
    kept, rows, cols, pids = accept_views(maps, min_visible)
    feats, seen = lift_mean(scene.n_points, scene.coords, lambda i: (scene.idmaps[kept[i]], scene.tables[kept[i]]), rows, cols, pids,
                            scene.tables.shape[2])
    lap("lift")
    vsem = scatter_mean(feats, scene.inds_reconstruct)
    vgeo = scatter_mean(scene.gauss[:, :6], scene.inds_reconstruct)
    vin = torch.cat([vsem, vgeo], 1)
    lap("scatter_mean")
    emb = student_forward(vin, scene.voxel_coords, state)
    lap("student")
    nbr = knn_c(scene.voxel_coords, k)
    lap("knn")
    w = affinity_weights(emb, nbr, sharpen)
    use_c = pool_impl == "c" or (pool_impl == "auto" and nbr.shape[0] * nbr.shape[1] > 2_000_000)
    pooled = pool_c(vin, nbr, w, n_iter, morton_order(scene.voxel_coords)) if use_c else pool(vin, nbr, w, n_iter)
    pts = pooled[torch.as_tensor(scene.inds_reconstruct)][:, :scene.tables.shape[2]]
    lap("pool")
    pred, logits = classify(pts, scene.text, scene.logit_scale, scene.coords[:, 1:4])
    i_, u_, t_ = intersection_and_union(pred, torch.as_tensor(scene.labels), scene.text.shape[0], (scene.ignore_label,))
    lap("classify")
    return dict(mappings=maps, kept=kept, lifted=feats, seen=seen, voxel_in=vin, embed=emb, knn=nbr,
                weights=w, pooled_voxels=pooled, scene_features=pts, pred=pred, logits=logits,
                inter=i_, union=u_, target=t_)
