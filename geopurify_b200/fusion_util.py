"""Drop-in mirror of the reference's models/utils/fusion_util.py mapper classes on the CUDA path.

Same constructor arguments, `compute_mapping` signature, return types (numpy int arrays) and
AssertionError behaviour as the reference (models/utils/fusion_util.py:36-147); the arithmetic
runs in gp_project_views / gp_splat_depth.  `compute_mapping_views` is the batched form (all
views of a scene in one launch, results left on the device).
"""
from __future__ import annotations

import math

import numpy as np
import torch

from . import _lib as L
from . import ops


def make_intrinsic(fx, fy, mx, my):
    intrinsic = np.eye(4)
    intrinsic[0][0], intrinsic[1][1], intrinsic[0][2], intrinsic[1][2] = fx, fy, mx, my
    return intrinsic


def adjust_intrinsic(intrinsic, intrinsic_image_dim, image_dim):
    if intrinsic_image_dim == image_dim:
        return intrinsic
    resize_width = int(math.floor(image_dim[1] * float(intrinsic_image_dim[0]) / float(intrinsic_image_dim[1])))
    intrinsic[0, 0] *= float(resize_width) / float(intrinsic_image_dim[0])
    intrinsic[1, 1] *= float(image_dim[1]) / float(intrinsic_image_dim[1])
    intrinsic[0, 2] *= float(image_dim[0] - 1) / float(intrinsic_image_dim[0] - 1)
    intrinsic[1, 2] *= float(image_dim[1] - 1) / float(intrinsic_image_dim[1] - 1)
    return intrinsic


def _coords_to_device(coords):
    c = coords if isinstance(coords, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(coords))
    assert c.dim() == 2 and c.shape[1] == 3, "[!] Shape error"
    if c.dtype not in (torch.float32, torch.float64):
        c = c.double()
    return c.cuda().contiguous()


def _depth_to_device(depth):
    if depth is None:
        return None
    d = depth if isinstance(depth, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(depth))
    if d.dtype not in (torch.float32, torch.float64):
        d = d.double()
    return d.cuda().contiguous()


class PointCloudToImageMapper(object):
    def __init__(self, image_dim, visibility_threshold=0.25, cut_bound=0, intrinsics=None):
        self.image_dim = image_dim
        self.vis_thres = visibility_threshold
        self.cut_bound = cut_bound
        self.intrinsics = np.array(intrinsics).copy()
        scale_x = self.image_dim[0] / (self.intrinsics[0, 2] * 2)
        scale_y = self.image_dim[1] / (self.intrinsics[1, 2] * 2)
        self.intrinsics[0, 0] *= scale_x
        self.intrinsics[1, 1] *= scale_y
        self.intrinsics[0, 2] = self.image_dim[0] / 2
        self.intrinsics[1, 2] = self.image_dim[1] / 2

    def _intr4(self, intrinsic=None):
        k = self.intrinsics if self.intrinsics is not None else intrinsic
        return np.array([k[0][0], k[1][1], k[0][2], k[1][2]], dtype=np.float64)

    def compute_mapping_views(self, world_to_camera_T, coords, depth=None):
        """All views at once.  world_to_camera_T [V,4,4] (each W2C^T, as the reference passes it),
        depth [V,H,W] | None | "splat".  Returns device tensors (pix [V,N] int32, counts [V] int32)."""
        L.require_cuda()
        W, H = int(self.image_dim[0]), int(self.image_dim[1])
        w2c = np.transpose(np.asarray(world_to_camera_T), (0, 2, 1))
        w2c_d, intr_d = ops.cameras_to_device(w2c, self._intr4())
        c = _coords_to_device(coords)
        if isinstance(depth, str):
            d = ops.splat_depth(c, w2c_d, intr_d, H, W, int(self.cut_bound))
        else:
            d = _depth_to_device(depth)
        return ops.project_views(c, w2c_d, intr_d, d, H, W, int(self.cut_bound), self.vis_thres)

    def compute_mapping(self, world_to_camera, coords, depth=None, intrinsic=None):
        """
        :param world_to_camera: 4 x 4 (W2C transposed, models/scene/camera.py:87)
        :param coords: N x 3 format
        :param depth: H x W format, None, or a string (z-buffer synthesis)
        :return: mapping, N x 3 format, (H,W,mask); weight N
        """
        coords_np = coords if isinstance(coords, np.ndarray) else np.asarray(coords)
        assert coords_np.ndim == 2 and coords_np.shape[1] == 3, "[!] Shape error"
        d = depth if (depth is None or isinstance(depth, str)) else np.asarray(depth)[None]
        pix, _ = self.compute_mapping_views(np.asarray(world_to_camera)[None], coords_np, d)
        mapping = ops.pix_to_mapping(pix[0], int(self.image_dim[0])).cpu().numpy().astype(int)
        # `weight` (fusion_util.py:118,145) is returned but never read by any caller (data_loader_ablation.py:249 drops
        # it): it is handed back as a lazy array that is only evaluated -- on the device -- if somebody looks at it.
        return mapping, LazyCenterWeight(self, np.asarray(world_to_camera), coords_np)


class LazyCenterWeight:
    """`weight = exp(-|round(pixel) - image centre| / 10)` of PointCloudToImageMapper.compute_mapping
    (models/utils/fusion_util.py:117-118, 145).  No caller reads it, so nothing is computed unless somebody looks:
    it behaves as a numpy array (`np.asarray(w)`, `w[i]`, `len(w)`, `w.shape`) and is evaluated on first use with
    the reference's own numpy formulas (bit-identical to its output; the golden tests compare it with atol = 0)."""

    def __init__(self, mapper, world_to_camera, coords):
        self._args = (mapper.intrinsics.copy(), tuple(mapper.image_dim), world_to_camera, coords)
        self._value = None
        self.shape = (coords.shape[0],)
        self.dtype = np.dtype(np.float64)

    def _eval(self):
        if self._value is None:
            k, dim, w2c_t, coords = self._args
            cn = np.concatenate([coords, np.ones([coords.shape[0], 1])], axis=1).T
            p = np.matmul(w2c_t.T, cn)
            with np.errstate(all="ignore"):
                p[0] = (p[0] * k[0][0]) / p[2] + k[0][2]
                p[1] = (p[1] * k[1][1]) / p[2] + k[1][2]
                pi = np.round(p).astype(int)
                cd = np.sqrt((pi[0] - dim[0] / 2) ** 2 + (pi[1] - dim[1] / 2) ** 2)
                self._value = np.exp(-cd / 10)
            self._args = None
        return self._value

    def __array__(self, dtype=None, copy=None):
        v = self._eval()
        return v if dtype is None else v.astype(dtype)

    def __len__(self):
        return self.shape[0]

    def __getitem__(self, i):
        return self._eval()[i]


class PointCloudToImageMappermatterport(object):
    def __init__(self, image_dim, visibility_threshold=0.25, cut_bound=0, intrinsics=None):
        self.image_dim = image_dim
        self.vis_thres = visibility_threshold
        self.cut_bound = cut_bound
        self.intrinsics = intrinsics

    def compute_mapping_views(self, camera_to_world, coords, depth, intrinsic):
        """camera_to_world [V,4,4]; intrinsic [V,3,3] or [3,3].  The 4x4 inverse stays on the host in the
        input dtype, exactly as the reference's np.linalg.inv (fusion_util.py:60)."""
        L.require_cuda()
        if self.intrinsics is not None:
            intrinsic = self.intrinsics
        W, H = int(self.image_dim[0]), int(self.image_dim[1])
        c2w = np.asarray(camera_to_world)
        w2c = np.stack([np.linalg.inv(c2w[v]) for v in range(c2w.shape[0])])
        k = np.asarray(intrinsic, dtype=np.float64)
        if k.ndim == 2:
            k = np.repeat(k[None], c2w.shape[0], 0)
        intr = np.stack([k[:, 0, 0], k[:, 1, 1], k[:, 0, 2], k[:, 1, 2]], 1)
        w2c_d, intr_d = ops.cameras_to_device(w2c, intr)
        return ops.project_views(_coords_to_device(coords), w2c_d, intr_d, _depth_to_device(depth), H, W,
                                 int(self.cut_bound), self.vis_thres)

    def compute_mapping(self, camera_to_world, coords, depth=None, intrinsic=None):
        coords_np = coords if isinstance(coords, np.ndarray) else np.asarray(coords)
        assert coords_np.ndim == 2 and coords_np.shape[1] == 3, "[!] Shape error"
        d = None if depth is None else np.asarray(depth)[None]
        pix, _ = self.compute_mapping_views(np.asarray(camera_to_world)[None], coords_np, d, intrinsic)
        return ops.pix_to_mapping(pix[0], int(self.image_dim[0])).cpu().numpy().astype(int)
