"""Synthetic ScanNet-/Matterport-shaped scenes for parity tests and bench.

Pure numpy, deterministic per seed.  Nothing here is on the timed path: it only
manufactures the inputs the reference would read from disk (`.pth` point clouds,
depth PNGs, poses) and the per-view output of the out-of-scope 2-D VLM.

Conventions follow the reference:
  * cameras: W2C = [R^T | t] as float32 (`models/utils/graphics_utils.py:42-53`);
    the mapper receives W2C^T (`models/scene/camera.py:87`).
  * intrinsics: ScanNet 577.870605 @ 640x480 adjusted to the working image size
    (`models/utils/mapping_util.py:37-52`, `models/utils/fusion_util.py:18-33`).
  * depth: z-buffer of the scene's own points, the reference's string-depth branch
    (`models/utils/fusion_util.py:126-130`), 3x3 min filter, 2 % invalid pixels.
  * voxels: floor(xyz / 0.02) - min, FNV-1-style key, `np.unique` order
    (`dataset/voxelizer.py:104-121`, `dataset/voxelization_utils.py:6-18, 78-96`).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

SCANNET_VAL_SIZES = dict(min=28_000, median=105_000, mean=115_000, max=302_000)


# --------------------------------------------------------------------------- intrinsics
def make_intrinsic(fx, fy, mx, my):
    k = np.eye(4)
    k[0, 0], k[1, 1], k[0, 2], k[1, 2] = fx, fy, mx, my
    return k


def adjust_intrinsic(intrinsic, intrinsic_image_dim, image_dim):
    """Same arithmetic as `models/utils/fusion_util.py:18-33` (returns a new array)."""
    intrinsic = np.array(intrinsic, dtype=np.float64).copy()
    if list(intrinsic_image_dim) == list(image_dim):
        return intrinsic
    resize_width = int(math.floor(image_dim[1] * float(intrinsic_image_dim[0]) / float(intrinsic_image_dim[1])))
    intrinsic[0, 0] *= float(resize_width) / float(intrinsic_image_dim[0])
    intrinsic[1, 1] *= float(image_dim[1]) / float(intrinsic_image_dim[1])
    intrinsic[0, 2] *= float(image_dim[0] - 1) / float(intrinsic_image_dim[0] - 1)
    intrinsic[1, 2] *= float(image_dim[1] - 1) / float(intrinsic_image_dim[1] - 1)
    return intrinsic


def scannet_intrinsics(image_dim=(320, 240)):
    k = make_intrinsic(577.870605, 577.870605, 319.5, 239.5)
    return adjust_intrinsic(k, [640, 480], list(image_dim))


def mapper_intrinsics(intrinsics, image_dim):
    """The rescale `PointCloudToImageMapper.__init__` applies (`fusion_util.py:90-96`)."""
    k = np.array(intrinsics, dtype=np.float64).copy()
    sx = image_dim[0] / (k[0, 2] * 2)
    sy = image_dim[1] / (k[1, 2] * 2)
    k[0, 0] *= sx
    k[1, 1] *= sy
    k[0, 2] = image_dim[0] / 2
    k[1, 2] = image_dim[1] / 2
    return k


# --------------------------------------------------------------------------- voxeliser
def fnv_hash_vec(arr):
    """FNV-1-style uint64 key, multiply then xor per coordinate (`voxelization_utils.py:6-18`)."""
    arr = np.asarray(arr).astype(np.uint64)
    h = np.full(arr.shape[0], np.uint64(14695981039346656037), dtype=np.uint64)
    with np.errstate(over="ignore"):
        for j in range(arr.shape[1]):
            h = h * np.uint64(1099511628211)
            h = np.bitwise_xor(h, arr[:, j])
    return h


def voxelize(coords, voxel_size=0.02):
    """Identity-augmentation `Voxelizer.voxelize` (`dataset/voxelizer.py:104-121`).

    Returns (voxel_coords [M,3] float64 integer-valued, inds [M], inds_reconstruct [N]).
    """
    scaled = np.floor(np.asarray(coords, dtype=np.float64) * (1.0 / voxel_size))
    scaled = np.floor(scaled - scaled.min(0))
    key = fnv_hash_vec(scaled)
    _, inds, inv = np.unique(key, return_index=True, return_inverse=True)
    return scaled[inds], inds, inv.astype(np.int64)


# --------------------------------------------------------------------------- geometry
def _sample_rect(rng, origin, eu, ev, n):
    a = rng.random((n, 1))
    b = rng.random((n, 1))
    return origin[None, :] + a * eu[None, :] + b * ev[None, :]


def _box_faces(lo, hi, inward, skip_bottom):
    """List of (origin, eu, ev, normal) rectangles of an axis-aligned box."""
    faces = []
    ext = hi - lo
    for ax in range(3):
        u, v = (ax + 1) % 3, (ax + 2) % 3
        for side in (0, 1):
            if skip_bottom and ax == 2 and side == 0:
                continue
            o = lo.copy()
            if side:
                o[ax] = hi[ax]
            eu = np.zeros(3)
            ev = np.zeros(3)
            eu[u] = ext[u]
            ev[v] = ext[v]
            nrm = np.zeros(3)
            nrm[ax] = 1.0 if side else -1.0
            if inward:
                nrm = -nrm
            faces.append((o, eu, ev, nrm))
    return faces


@dataclass
class Scene:
    coords: np.ndarray            # [N,3] float32 world xyz
    gauss: np.ndarray             # [N,6] float32 rgb(0..1) | normal
    labels: np.ndarray            # [N] int64, `ignore_label` on 2 %
    w2c: np.ndarray               # [V,4,4] float32
    intrinsics: np.ndarray        # [4,4] float64, as handed to the mapper ctor
    image_dim: tuple              # (W, H)
    depth: np.ndarray             # [V,H,W] float64 metres
    idmaps: np.ndarray            # [V,H,W] uint8 region id per pixel
    tables: np.ndarray            # [V,Q,D] float32 region embeddings (rows >= nq[v] are zero)
    nq: np.ndarray                # [V] regions per view
    text: np.ndarray              # [C,D] float32
    logit_scale: float
    voxel_coords: np.ndarray      # [M,3] float32 integer-valued
    inds_reconstruct: np.ndarray  # [N] int64 point -> voxel
    vis_thres: float
    cut: int
    ignore_label: int = 255
    meta: dict = field(default_factory=dict)

    @property
    def n_points(self):
        return self.coords.shape[0]

    @property
    def n_views(self):
        return self.w2c.shape[0]

    def dense_map(self, v, dtype=np.float32):
        """[H,W,D] channels-last feature map of view v (table lookup of the id map)."""
        return self.tables[v][self.idmaps[v]].astype(dtype)


def make_scene(n_points=150_000, n_views=50, image_dim=(320, 240), feat_dim=512, n_classes=20,
               seed=5557, room=(5.0, 4.0, 2.7), n_boxes=8, voxel_size=0.02, vis_thres=0.05, cut=10,
               q_range=(20, 120), ignore_frac=0.02, invalid_depth_frac=0.02, noise=0.005,
               ignore_label=255):
    rng = np.random.default_rng(seed)
    W, H = image_dim
    room = np.asarray(room, dtype=np.float64)

    # ---- surfaces: room shell (normals inward) + boxes on the floor
    faces = _box_faces(np.zeros(3), room, inward=True, skip_bottom=False)
    boxes = []
    for _ in range(n_boxes):
        size = np.array([rng.uniform(0.4, 1.6), rng.uniform(0.4, 1.2), rng.uniform(0.3, 1.4)])
        size = np.minimum(size, room * 0.45)
        lo = np.array([rng.uniform(0.05, room[0] - size[0] - 0.05),
                       rng.uniform(0.05, room[1] - size[1] - 0.05), 0.0])
        boxes.append((lo, lo + size))
        faces += _box_faces(lo, lo + size, inward=False, skip_bottom=True)
    areas = np.array([np.linalg.norm(eu) * np.linalg.norm(ev) for _, eu, ev, _ in faces])
    counts = rng.multinomial(n_points, areas / areas.sum())
    pts, nrms, surf = [], [], []
    for fi, ((o, eu, ev, nrm), c) in enumerate(zip(faces, counts)):
        if c == 0:
            continue
        pts.append(_sample_rect(rng, o, eu, ev, c))
        nrms.append(np.repeat(nrm[None, :], c, 0))
        surf.append(np.full(c, fi))
    pts = np.concatenate(pts) + rng.normal(0.0, noise, (n_points, 3))
    nrms = np.concatenate(nrms)
    surf = np.concatenate(surf)
    perm = rng.permutation(n_points)
    coords = pts[perm].astype(np.float32)
    nrms = nrms[perm].astype(np.float32)
    surf = surf[perm]
    rgb = rng.random((n_points, 3)).astype(np.float32)
    gauss = np.concatenate([rgb, nrms], 1).astype(np.float32)

    # ---- labels: one hidden class per surface
    face_class = rng.integers(0, n_classes, len(faces))
    true_class = face_class[surf].astype(np.int64)
    labels = true_class.copy()
    labels[rng.random(n_points) < ignore_frac] = ignore_label

    # ---- text embeddings
    text = rng.normal(size=(n_classes, feat_dim))
    text /= np.linalg.norm(text, axis=1, keepdims=True)
    text = text.astype(np.float32)

    # ---- cameras
    intr = scannet_intrinsics(image_dim) if tuple(image_dim) == (320, 240) else \
        make_intrinsic(577.870605 * W / 640.0, 577.870605 * W / 640.0, (W - 1) / 2.0, (H - 1) / 2.0)
    kmap = mapper_intrinsics(intr, image_dim)
    w2c = np.zeros((n_views, 4, 4), dtype=np.float32)
    for v in range(n_views):
        for _ in range(100):
            c = np.array([rng.uniform(0.3, room[0] - 0.3), rng.uniform(0.3, room[1] - 0.3),
                          rng.uniform(1.2, min(1.8, room[2] - 0.2))])
            if not any((c[:2] > lo[:2] - 0.1).all() and (c[:2] < hi[:2] + 0.1).all() and c[2] < hi[2] + 0.1
                       for lo, hi in boxes):
                break
        yaw = rng.uniform(0, 2 * np.pi)
        pitch = np.deg2rad(rng.uniform(-30.0, 10.0))
        f = np.array([np.cos(yaw) * np.cos(pitch), np.sin(yaw) * np.cos(pitch), np.sin(pitch)])
        right = np.array([f[1], -f[0], 0.0])
        right /= np.linalg.norm(right)
        down = np.cross(f, right)
        r_c2w = np.stack([right, down, f], 1)
        m = np.eye(4)
        m[:3, :3] = r_c2w.T
        m[:3, 3] = -r_c2w.T @ c
        w2c[v] = m.astype(np.float32)

    # ---- depth (z-buffer of own points, min filter, invalid pixels) and 2-D regions
    c64 = coords.astype(np.float64)
    homo = np.concatenate([c64, np.ones((n_points, 1))], 1).T
    depth = np.zeros((n_views, H, W), dtype=np.float64)
    qmax = q_range[1]
    idmaps = np.zeros((n_views, H, W), dtype=np.uint8)
    tables = np.zeros((n_views, qmax, feat_dim), dtype=np.float32)
    nq = np.zeros(n_views, dtype=np.int64)
    yy, xx = np.mgrid[0:H, 0:W]
    for v in range(n_views):
        p = w2c[v].astype(np.float64) @ homo
        u = np.round(p[0] * kmap[0, 0] / p[2] + kmap[0, 2])
        w_ = np.round(p[1] * kmap[1, 1] / p[2] + kmap[1, 2])
        ok = (u >= cut) & (w_ >= cut) & (u < W - cut) & (w_ < H - cut) & (p[2] > 0.2)
        ui, vi, z = u[ok].astype(np.int64), w_[ok].astype(np.int64), p[2][ok]
        d = np.full(H * W, 999999.0)
        np.minimum.at(d, vi * W + ui, z)
        d = d.reshape(H, W)
        # 3x3 min filter over non-empty pixels closes splat holes
        pad = np.pad(d, 1, constant_values=999999.0)
        stack = np.stack([pad[i:i + H, j:j + W] for i in range(3) for j in range(3)])
        d = stack.min(0)
        d[d >= 999999.0] = 0.0
        d[rng.random((H, W)) < invalid_depth_frac] = 0.0
        depth[v] = d

        # Voronoi partition into q regions; embedding = text[majority class] + noise
        q = int(rng.integers(q_range[0], q_range[1] + 1))
        nq[v] = q
        seeds = np.stack([rng.uniform(0, H, q), rng.uniform(0, W, q)], 1)
        d2 = (yy[None] - seeds[:, 0, None, None]) ** 2 + (xx[None] - seeds[:, 1, None, None]) ** 2
        idm = d2.argmin(0).astype(np.uint8)
        idmaps[v] = idm
        vis = ok & (np.abs(d[np.clip(w_, 0, H - 1).astype(np.int64), np.clip(u, 0, W - 1).astype(np.int64)] - p[2])
                    <= vis_thres * d[np.clip(w_, 0, H - 1).astype(np.int64), np.clip(u, 0, W - 1).astype(np.int64)])
        reg = idm[w_[vis].astype(np.int64), u[vis].astype(np.int64)]
        votes = np.zeros((q, n_classes), dtype=np.int64)
        np.add.at(votes, (reg, true_class[vis]), 1)
        rc = np.where(votes.sum(1) > 0, votes.argmax(1), rng.integers(0, n_classes, q))
        emb = 0.55 * text[rc].astype(np.float64) + rng.normal(size=(q, feat_dim)) / math.sqrt(feat_dim)
        emb /= np.linalg.norm(emb, axis=1, keepdims=True)
        tables[v, :q] = emb.astype(np.float32)

    vox, _, inv = voxelize(coords, voxel_size)
    return Scene(coords=coords, gauss=gauss, labels=labels, w2c=w2c, intrinsics=intr, image_dim=(W, H),
                 depth=depth, idmaps=idmaps, tables=tables, nq=nq, text=text, logit_scale=100.0,
                 voxel_coords=vox.astype(np.float32), inds_reconstruct=inv, vis_thres=vis_thres, cut=cut,
                 ignore_label=ignore_label,
                 meta=dict(seed=seed, room=tuple(room), n_boxes=n_boxes, voxel_size=voxel_size))


# --------------------------------------------------------------------------- student weights
STUDENT_LAYERS = (  # name, kernel volume, cin, cout   (models/affinity_module.py:55-66, 266-270)
    ("input_layer.0", 27, None, None),
)


def make_student_state(input_dim=518, hidden_dim=512, embed_dim=128, seed=5557, dtype=np.float32):
    """Random-init student in MinkowskiEngine's state-dict naming."""
    rng = np.random.default_rng(seed)
    sd = {}

    def conv(name, kv, cin, cout):
        w = rng.normal(0.0, math.sqrt(2.0 / (kv * cin)), (kv, cin, cout)).astype(dtype)
        sd[name + ".kernel"] = w if kv > 1 else w[0]

    def bn(name, c):
        sd[name + ".bn.weight"] = rng.uniform(0.8, 1.2, c).astype(dtype)
        sd[name + ".bn.bias"] = rng.normal(0, 0.05, c).astype(dtype)
        sd[name + ".bn.running_mean"] = rng.normal(0, 0.05, c).astype(dtype)
        sd[name + ".bn.running_var"] = rng.uniform(0.8, 1.2, c).astype(dtype)

    conv("input_layer.0", 27, input_dim, hidden_dim)
    bn("input_layer.1", hidden_dim)
    for b in range(4):
        conv(f"res_blocks.{b}.conv1", 27, hidden_dim, hidden_dim)
        bn(f"res_blocks.{b}.norm1", hidden_dim)
        conv(f"res_blocks.{b}.conv2", 27, hidden_dim, hidden_dim)
        bn(f"res_blocks.{b}.norm2", hidden_dim)
    conv("output_layer", 1, hidden_dim, embed_dim)
    return sd


def scannet_val_size(rng):
    """Draw N from a log-normal fitted to `dataset/scannet_val_metrics.tsv` (min 28k, median 105k, max 302k)."""
    n = int(np.exp(rng.normal(math.log(105_000), 0.42)))
    return int(np.clip(n, 28_000, 302_000))


def make_mask_predictions(n_views, n_queries, image_hw, text, feat_dim=512, seed=7):
    """Seeded stand-in for the X-Decoder head outputs of models/affinity_module.py:519-537, per view:
    `pred_masks` [V,Q,H,W] float32 logits (blobs; about a fifth of the image stays below 0 for every query = "no
    mask", and some queries never win a pixel), `pred_logits` [V,Q,C+1] float32, `mask_embed` [V,Q,D] float32
    (class text embedding + noise).  Everything is computed in float64 and rounded once, so the values are
    reproducible from the seed alone."""
    rng = np.random.default_rng(seed)
    H, W = image_hw
    C = text.shape[0]
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float64)
    masks = np.empty((n_views, n_queries, H, W), dtype=np.float32)
    logits = np.empty((n_views, n_queries, C + 1), dtype=np.float32)
    embed = np.empty((n_views, n_queries, feat_dim), dtype=np.float32)
    for v in range(n_views):
        for q in range(n_queries):
            cy, cx = rng.uniform(0, H), rng.uniform(0, W)
            rad = rng.uniform(0.12, 0.3) * max(H, W)
            amp = rng.uniform(2.0, 6.0)
            masks[v, q] = (amp * (1.0 - ((yy - cy) ** 2 + (xx - cx) ** 2) / rad ** 2)).astype(np.float32)
            cls = int(rng.integers(0, C))
            row = rng.normal(0.0, 1.0, C + 1)
            row[cls] += rng.uniform(1.0, 4.0)
            logits[v, q] = row.astype(np.float32)
            embed[v, q] = (text[cls].astype(np.float64) + rng.normal(0.0, 0.05, feat_dim)).astype(np.float32)
    return masks, logits, embed


def make_mining_inputs(n=1500, d=64, seed=11, n_clusters=10):
    """Seeded inputs for the training-side sampling kernels (F4): random points in a room and clustered features."""
    rng = np.random.default_rng(seed)
    xyz = (rng.random((n, 3)) * np.array([4.0, 3.0, 2.5])).astype(np.float32)
    proto = rng.standard_normal((n_clusters, d)).astype(np.float32)
    feat = (proto[rng.integers(0, n_clusters, n)] + 0.7 * rng.standard_normal((n, d))).astype(np.float32)
    return xyz, feat
