"""Generate golden vectors for SURVEY.md §8 row A1/A2 from the REFERENCE'S OWN code.

Run in the build container only (needs /root/reference, which does not travel to the GPU box):

    python tests/golden/make_golden.py

Imports `models/utils/fusion_util.py` from the reference tree by path (numpy only), runs
`PointCloudToImageMapper.compute_mapping` and `PointCloudToImageMappermatterport.compute_mapping`
on small seeded inputs and writes inputs + outputs to `tests/golden/project_*.npz`.
Nothing is copied from the reference; only its outputs are stored.
"""
import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from geopurify_b200 import synth  # noqa: E402

REF = "/root/reference/models/utils/fusion_util.py"


def load_ref():
    spec = importlib.util.spec_from_file_location("ref_fusion_util", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ref = load_ref()
    cases = {}

    # ---- case 1: ScanNet-style mapper on a small synthetic room, float64 coords + float64 depth
    sc = synth.make_scene(n_points=6000, n_views=4, image_dim=(80, 60), seed=11, room=(2.5, 2.0, 2.4),
                          n_boxes=3, cut=3, q_range=(4, 9))
    W, H = sc.image_dim
    mapper = ref.PointCloudToImageMapper([W, H], sc.vis_thres, sc.cut, sc.intrinsics)
    c64 = sc.coords.astype(np.float64)
    outs, weights = [], []
    for v in range(sc.n_views):
        m, w = mapper.compute_mapping(sc.w2c[v].T.copy(), c64, sc.depth[v])
        outs.append(m)
        weights.append(w)
    cases["scannet_f64"] = dict(coords=c64, w2c_T=np.stack([sc.w2c[v].T.copy() for v in range(sc.n_views)]),
                                depth=sc.depth, intrinsics=sc.intrinsics, image_dim=np.array([W, H]),
                                vis_thres=sc.vis_thres, cut=sc.cut, mapping=np.stack(outs), weight=np.stack(weights))

    # ---- case 2: same, depth=None branch (front test only) and float32 coords
    outs = [mapper.compute_mapping(sc.w2c[v].T.copy(), sc.coords, None)[0] for v in range(sc.n_views)]
    cases["scannet_nodepth_f32coords"] = dict(coords=sc.coords, w2c_T=cases["scannet_f64"]["w2c_T"],
                                              intrinsics=sc.intrinsics, image_dim=np.array([W, H]),
                                              vis_thres=sc.vis_thres, cut=sc.cut, mapping=np.stack(outs))

    # ---- case 3: string-depth branch (z-buffer synthesis, fusion_util.py:126-130)
    outs = [mapper.compute_mapping(sc.w2c[v].T.copy(), c64, "splat")[0] for v in range(2)]
    cases["scannet_splat"] = dict(coords=c64, w2c_T=cases["scannet_f64"]["w2c_T"][:2], intrinsics=sc.intrinsics,
                                  image_dim=np.array([W, H]), vis_thres=sc.vis_thres, cut=sc.cut,
                                  mapping=np.stack(outs))

    # ---- case 4: float32 depth (numpy keeps vis_thres * depth in float32)
    d32 = sc.depth.astype(np.float32)
    outs = [mapper.compute_mapping(sc.w2c[v].T.copy(), c64, d32[v])[0] for v in range(sc.n_views)]
    cases["scannet_f32depth"] = dict(coords=c64, w2c_T=cases["scannet_f64"]["w2c_T"], depth=d32,
                                     intrinsics=sc.intrinsics, image_dim=np.array([W, H]),
                                     vis_thres=sc.vis_thres, cut=sc.cut, mapping=np.stack(outs))

    # ---- case 5: edge cases -- exact half-integer pixels (round-half-even), z == 0, z < 0, borders, cut = 0
    Wd, Hd = 16, 12
    intr = synth.make_intrinsic(1.0, 1.0, Wd / 2, Hd / 2)   # rescale in __init__ is then the identity
    mapper_e = ref.PointCloudToImageMapper([Wd, Hd], 0.25, 0, intr)
    xs = np.array([-8.5, -8.0, -7.5, -0.5, 0.0, 0.5, 1.5, 2.5, 6.5, 7.0, 7.49999, 7.5, 8.0, 1e300, -1e300, np.nan])
    ys = np.array([-6.5, -6.0, -5.5, -0.5, 0.5, 1.5, 4.5, 5.0, 5.49999, 5.5, 6.0])
    zs = np.array([1.0, 0.0, -1.0, 2.0, 1e-320])
    g = np.array([(x, y, z) for x in xs for y in ys for z in zs], dtype=np.float64)
    eye_T = np.eye(4, dtype=np.float32)
    depth_e = np.ones((Hd, Wd)) * 1.0
    depth_e[::2, ::3] = 0.0
    depth_e[1::4, 1::5] = 2.0
    with np.errstate(all="ignore"):
        m_d = mapper_e.compute_mapping(eye_T, g, depth_e)[0]
        m_n = mapper_e.compute_mapping(eye_T, g, None)[0]
    cases["edge"] = dict(coords=g, w2c_T=eye_T[None], depth=depth_e[None], intrinsics=intr,
                         image_dim=np.array([Wd, Hd]), vis_thres=0.25, cut=0, mapping=m_d[None],
                         mapping_nodepth=m_n[None])

    # ---- case 6: Matterport mapper: camera_to_world (float32) inverted inside, 3x3 intrinsics
    scm = synth.make_scene(n_points=5000, n_views=3, image_dim=(64, 52), seed=12, room=(3.0, 2.5, 2.5),
                           n_boxes=2, cut=2, vis_thres=0.02, q_range=(4, 9))
    Wm, Hm = scm.image_dim
    k3 = synth.mapper_intrinsics(scm.intrinsics, scm.image_dim)[:3, :3].copy()
    mapper_m = ref.PointCloudToImageMappermatterport([Wm, Hm], 0.02, 2)
    c2w = np.stack([np.linalg.inv(scm.w2c[v].astype(np.float64)).astype(np.float32) for v in range(3)])
    cm = scm.coords.astype(np.float64)
    outs = [mapper_m.compute_mapping(c2w[v], cm, scm.depth[v], k3) for v in range(3)]
    cases["matterport"] = dict(coords=cm, c2w=c2w, depth=scm.depth, intrinsics=k3, image_dim=np.array([Wm, Hm]),
                               vis_thres=0.02, cut=2, mapping=np.stack(outs))

    for name, d in cases.items():
        path = os.path.join(HERE, f"project_{name}.npz")
        np.savez_compressed(path, **d)
        print(name, {k: getattr(v, "shape", v) for k, v in d.items()}, os.path.getsize(path))


if __name__ == "__main__":
    main()


def intrinsics_golden():
    """make_intrinsic / adjust_intrinsic of the reference (models/utils/fusion_util.py:7-33) on four camera models
    -> tests/golden/intrinsics.json (called from main below when run as a script)."""
    import json
    ref = load_ref()
    cases = []
    for fx, fy, mx, my, src, dst in [(577.870605, 577.870605, 319.5, 239.5, [640, 480], [320, 240]),
                                     (1170.18, 1170.18, 647.75, 483.75, [1296, 968], [648, 484]),
                                     (1075.4, 1076.1, 629.7, 522.3, [1280, 1024], [640, 512]),
                                     (500.0, 510.0, 100.0, 80.0, [200, 160], [200, 160])]:
        k = ref.make_intrinsic(fx, fy, mx, my)
        a = ref.adjust_intrinsic(k.copy(), src, dst)
        cases.append(dict(args=[fx, fy, mx, my], src=src, dst=dst, make=k.tolist(), adjust=np.asarray(a).tolist()))
    json.dump(cases, open(os.path.join(HERE, "intrinsics.json"), "w"), indent=1)


if __name__ == "__main__":
    intrinsics_golden()
