"""Mirror of the reference's `Voxelizer` (dataset/voxelizer.py:41-132) on the CUDA path -- SURVEY.md §8 row F1.

Only the inference configuration is implemented: no clipping (`clip_bound=None`) and either no augmentation
(`use_augmentation=False`, every shipped eval config) or a caller-supplied rigid transformation.  The quantisation
itself -- floor, min shift, FNV-1-style key, np.unique order, first-point representative -- runs in `gp_voxelize`.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L
from . import ops


class Voxelizer:
    def __init__(self, voxel_size=1, clip_bound=None, use_augmentation=False, scale_augmentation_bound=None,
                 rotation_augmentation_bound=None, translation_augmentation_ratio_bound=None, ignore_label=255):
        if clip_bound is not None:
            raise NotImplementedError("geopurify_b200.Voxelizer: clip_bound is a training-time option (out of scope)")
        if use_augmentation:
            raise NotImplementedError("geopurify_b200.Voxelizer: random augmentation is out of scope; pass the "
                                      "matrix you want through voxelize(..., rigid_transformation=M)")
        self.voxel_size = voxel_size
        self.ignore_label = ignore_label

    def get_transformation_matrix(self):
        """(M_v, M_r) of dataset/voxelizer.py:59-79 without augmentation: M_v = diag(1/voxel_size), M_r = I."""
        m_v, m_r = np.eye(4), np.eye(4)
        np.fill_diagonal(m_v[:3, :3], 1 / self.voxel_size)
        return m_v, m_r

    def voxelize_device(self, coords, rigid_transformation=None):
        """coords: CUDA tensor [N,3] float32/float64.  Returns (vox [M,3] int32, inds [M] int32,
        inds_reconstruct [N] int32, extent) on the device."""
        L.require_cuda()
        m = self.get_transformation_matrix()[0] if rigid_transformation is None else np.asarray(rigid_transformation)
        return ops.voxelize(coords.contiguous(), np.ascontiguousarray(m[:3, :4], dtype=np.float64))

    def voxelize(self, coords, feats, labels, center=None, link=None, return_ind=False, rigid_transformation=None):
        """Reference signature and return values (numpy): coords_aug [M,3] float64 integer-valued, feats[inds],
        labels[inds], inds_reconstruct [N] int64 (, inds | link[inds])."""
        assert coords.shape[1] == 3 and coords.shape[0] == feats.shape[0] and coords.shape[0]
        c = torch.as_tensor(np.ascontiguousarray(coords))
        if c.dtype not in (torch.float32, torch.float64):
            c = c.double()
        vox, inds, rec, _ = self.voxelize_device(c.cuda(), rigid_transformation)
        inds = inds.cpu().numpy().astype(np.int64)
        coords_aug = vox.cpu().numpy().astype(np.float64)
        feats, labels = feats[inds], labels[inds]
        rec = rec.cpu().numpy().astype(np.int64)
        if return_ind:
            return coords_aug, feats, labels, rec, inds
        if link is not None:
            return coords_aug, feats, labels, rec, link[inds]
        return coords_aug, feats, labels, rec
