"""BASELINE.json configs[4]: kNN + pooling scaling sweep on one B200 (device-resident), one JSON line per point.

    python scripts/sweep_pool_knn.py [--points 50000,100000,...] [--K 16,32,64,96] [--T 1,3,5,10,19]

`frac` = algorithmic bytes / time against the measured HBM peak: kNN m*16 + m*K*4; pooling T*(2*m*D*4 + m*K*4 + m*E*4)
(SURVEY.md §8(d)).  Geometry: the synthetic room scaled so that the ScanNet point density is kept."""
import argparse, json, math, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from geopurify_b200 import synth, ops

ap = argparse.ArgumentParser()
ap.add_argument("--points", default="50000,100000,200000,500000,1000000,2000000")
ap.add_argument("--K", default="16,32,64,96")
ap.add_argument("--T", default="1,3,5,10,19")
a = ap.parse_args()
peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))) \
    if os.path.exists("MEASURED_PEAKS.json") else {"hbm_gbs": 6650.0}
HBM = float(peaks["hbm_gbs"])


def timeit(fn, reps=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


for n in [int(x) for x in a.points.split(",")]:
    s = math.sqrt(n / 150_000.0)
    sc = synth.make_scene(n_points=n, n_views=1, image_dim=(32, 24), feat_dim=8, n_classes=3, seed=5557,
                          room=(5.0 * s, 4.0 * s, 2.7), n_boxes=max(2, int(8 * s * s)))
    vox = torch.as_tensor(sc.voxel_coords.astype(np.int32)).cuda()
    m = vox.shape[0]
    idx = ops.VoxelIndex(vox)
    emb = torch.nn.functional.normalize(torch.randn(m, 128, device="cuda"), dim=1)
    f = torch.randn(m, 544, device="cuda") * 0.05
    for K in [int(x) for x in a.K.split(",")]:
        t_knn, (_, nbr) = timeit(lambda: idx.knn(K, want_orig=False))
        w = ops.affinity_softmax(emb, nbr, 20.0)
        knn_bytes = m * 16 + m * K * 4
        line = {"stage": "knn", "points": n, "voxels": m, "K": K, "ms": round(t_knn, 4),
                "GBps": round(knn_bytes / t_knn / 1e6, 1), "frac": round(knn_bytes / t_knn / 1e6 / HBM, 4)}
        print(json.dumps(line), flush=True)
        for T in [int(x) for x in a.T.split(",")]:
            t_pool, _ = timeit(lambda: ops.pool(f, 512, nbr, w, T))
            pb = T * (2 * m * 512 * 4 + m * K * 4 + m * 128 * 4)
            line = {"stage": "pool", "points": n, "voxels": m, "K": K, "T": T, "ms": round(t_pool, 4),
                    "ms_per_application": round(t_pool / T, 4), "GBps": round(pb / t_pool / 1e6, 1),
                    "frac": round(pb / t_pool / 1e6 / HBM, 4)}
            print(json.dumps(line), flush=True)
    del idx, emb, f
    torch.cuda.empty_cache()
