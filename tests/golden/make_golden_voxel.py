"""Golden vectors for SURVEY.md §8 row F1 from the REFERENCE'S OWN voxeliser.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_voxel.py

`dataset/voxelization_utils.py:3` and `dataset/voxelizer.py:37` use `collections.Sequence` / `collections.Iterable`,
aliases Python 3.10 removed; the two names are re-bound to their `collections.abc` homes before the import and the
reference code then runs unmodified (`Voxelizer.voxelize` -> `sparse_quantize` -> `fnv_hash_vec` -> `np.unique`).
Only inputs and outputs are stored.
"""
import collections
import collections.abc
import os
import sys

import numpy as np

collections.Sequence = collections.abc.Sequence
collections.Iterable = collections.abc.Iterable
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, "/root/reference")
from dataset.voxelizer import Voxelizer  # noqa: E402
from geopurify_b200 import synth  # noqa: E402


def run(coords, voxel_size):
    v = Voxelizer(voxel_size=voxel_size, clip_bound=None, use_augmentation=False)
    n = coords.shape[0]
    feats = np.arange(n * 3, dtype=np.float64).reshape(n, 3)
    labels = np.arange(n, dtype=np.int64)
    c, f, l, rec, inds = v.voxelize(coords, feats, labels, return_ind=True)
    return dict(coords=coords, voxel_size=voxel_size, voxel_coords=c, inds=inds, inds_reconstruct=rec)


def main():
    rng = np.random.default_rng(5557)
    sc = synth.make_scene(n_points=6000, n_views=2, image_dim=(80, 60), seed=11, room=(2.5, 2.0, 2.4), n_boxes=3,
                          cut=3, q_range=(4, 9))
    cases = {
        "surface_f32": run(sc.coords.astype(np.float32), 0.02),                     # the loader's case: 2 cm voxels
        "cloud_f64_negative": run((rng.random((5000, 3)) - 0.5) * 3.0, 0.1),        # heavy collisions, negative coords
        "single_point": run(np.array([[0.3, -1.2, 2.0]]), 0.02),
        "duplicates": run(np.repeat(rng.random((40, 3)), 5, axis=0)[rng.permutation(200)], 0.05),
        "on_cell_faces": run(np.round(rng.random((500, 3)) * 20) * 0.05, 0.05),     # coordinates on voxel boundaries
    }
    for name, c in cases.items():
        np.savez_compressed(os.path.join(HERE, f"voxel_{name}.npz"), **c)
        print(name, c["coords"].shape, "->", c["voxel_coords"].shape)


if __name__ == "__main__":
    main()
