"""Whole-scene parity at the shapes BASELINE.json names (scaled to sizes the CPU oracle finishes in seconds):
the reference's hard-coded K=96 / T=19 / D=512 / hidden 512 (configs[0,1]), 200 classes (configs[2]),
Matterport-style tau=0.02 + 160 classes (configs[3]) and the K / T sweep values (configs[4])."""
import numpy as np
import pytest
import torch

from geopurify_b200 import synth

pytestmark = pytest.mark.gpu

CASES = {
    # name: (n_points, image_dim, n_classes, vis_thres, hidden, K, T)
    "reference_constants": (9000, (80, 60), 20, 0.05, 512, 96, 19),
    "scannet200_classes": (9000, (80, 60), 200, 0.05, 128, 32, 1),
    "matterport_like": (10000, (160, 128), 160, 0.02, 128, 64, 10),
    "sweep_k16_t5": (9000, (80, 60), 20, 0.05, 128, 16, 5),
}


@pytest.mark.parametrize("name", list(CASES))
def test_whole_scene_config(name):
    from geopurify_b200 import _lib
    from geopurify_b200.affinity_module import SonataXAffinityTrainer, scene_inputs_from_synth
    from geopurify_b200.student import AffinityPredictor, load_numpy_state
    from geopurify_b200.validation import mean_iou
    from oracle import oracle
    _lib.require_cuda()
    n, dim, C, tau, hidden, K, T = CASES[name]
    sc = synth.make_scene(n_points=n, n_views=6, image_dim=dim, seed=21, room=(2.0, 1.6, 2.0), n_boxes=2, cut=2,
                          q_range=(4, 12), feat_dim=512, n_classes=C, vis_thres=tau)
    state = synth.make_student_state(input_dim=518, hidden_dim=hidden, embed_dim=128, seed=21)
    student = load_numpy_state(AffinityPredictor(518, 128, hidden), state)
    model = SonataXAffinityTrainer(text_features=sc.text, logit_scale=sc.logit_scale, student=student,
                                   feature_dim=512, K=K, n_pool=T, min_visible=50)
    s = scene_inputs_from_synth(sc, feat_dtype=torch.float32)
    ign = torch.tensor([sc.ignore_label], dtype=torch.int64, device="cuda")
    out = model.run_scene(s, ignore=ign, keep_intermediates=True)
    torch.cuda.synchronize()
    ref = oracle.evaluate_scene(sc, state, k=K, n_iter=T, min_visible=50)

    order = model.last["index"].order.long().cpu()
    assert (model.last_seen.cpu().numpy().astype(bool) == ref["seen"].numpy()).all()
    assert (model.last["knn_orig"].cpu().numpy().astype(np.int64) == ref["knn"]).all(), "kNN indices differ"
    e_emb = (model.last["embed"].cpu() - ref["embed"][order]).abs().max().item()
    e_pool = (model.last["pooled"].cpu() - ref["pooled_voxels"][order][:, :512]).abs().max().item()
    assert e_emb < 2.5e-5, f"embedding max-abs {e_emb}"
    assert e_pool < 1e-3, f"pooled max-abs {e_pool} (north star bound 1e-3)"

    pred = out["pred"].cpu()
    ref_pred = ref["pred"].clone()
    ref_pred[torch.as_tensor(sc.labels) == sc.ignore_label] = sc.ignore_label
    mism = int((pred != ref_pred).sum())
    print(f"{name}: M={order.numel()} emb {e_emb:.2e} pooled {e_pool:.2e} label mismatches {mism}")
    assert mism == 0, f"{mism} labels differ from the oracle (north star: bit-exact)"
    h = out["hist"].cpu().numpy()
    assert (h[0] == ref["inter"].numpy()).all() and (h[1] == ref["union"].numpy()).all()
    assert (h[2] == ref["target"].numpy()).all()
    assert mean_iou(h[0], h[1], h[2]) == oracle.miou(ref["inter"].numpy(), ref["union"].numpy(),
                                                     ref["target"].numpy())


def test_reference_tuple_surface_consensus_and_mean(small_state):
    """evaluate_scene on the reference's positional 20-tuple (CPU tensors in, dict out) for both teachers' lifting."""
    from geopurify_b200.affinity_module import SonataXAffinityTrainer
    from geopurify_b200.student import AffinityPredictor, load_numpy_state
    from geopurify_b200 import ops
    from oracle import oracle
    sc = synth.make_scene(n_points=8000, n_views=5, image_dim=(80, 60), seed=9, room=(2.0, 1.6, 2.0), n_boxes=2,
                          cut=2, q_range=(4, 12), feat_dim=128, n_classes=7)
    W, H = sc.image_dim
    kmap = oracle.mapper_intrinsics(sc.intrinsics, sc.image_dim)
    maps = [oracle.project_view(sc.w2c[v].T.copy(), sc.coords.astype(np.float64), sc.depth[v], kmap, (W, H),
                                sc.vis_thres, sc.cut)[0] for v in range(sc.n_views)]
    kept = [v for v in range(sc.n_views) if maps[v][:, 2].sum() >= 50]
    n = sc.n_points
    x = torch.cat([torch.as_tensor(maps[v][maps[v][:, 2] == 1, 0]) for v in kept])
    y = torch.cat([torch.as_tensor(maps[v][maps[v][:, 2] == 1, 1]) for v in kept])
    mask2d = torch.cat([torch.stack([torch.full((n,), i), torch.as_tensor(maps[v][:, 2])], 1)
                        for i, v in enumerate(kept)])
    batch = [None] * 20
    batch[0] = torch.as_tensor(sc.coords)
    batch[1] = torch.as_tensor(sc.voxel_coords).float()
    batch[2] = torch.as_tensor(sc.inds_reconstruct)
    batch[3] = torch.as_tensor(sc.labels)
    batch[12], batch[13], batch[14] = x, y, mask2d
    batch[19] = torch.as_tensor(sc.gauss)
    idm = torch.as_tensor(sc.idmaps[kept]).cuda()
    tab = torch.as_tensor(sc.tables[kept]).cuda()
    student = load_numpy_state(AffinityPredictor(134, 128, 128), small_state)
    common = dict(text_features=sc.text, logit_scale=sc.logit_scale, student=student, feature_dim=128, K=16, n_pool=3)
    mean_model = SonataXAffinityTrainer(view_features=lambda b: ops.expand_idmap(idm, tab, torch.float32), **common)
    cons_model = SonataXAffinityTrainer(mask_features=lambda b: (idm, tab), **common)
    assert mean_model.use_lseg and not cons_model.use_lseg
    r_mean = mean_model.evaluate_scene(tuple(batch))
    r_cons = cons_model.evaluate_scene(tuple(batch))
    # any integer voxel origin (ME-style negative coordinates): same features, nothing out of bounds
    shifted = list(batch)
    shifted[1] = batch[1] - 37.0
    assert torch.equal(mean_model.evaluate_scene(tuple(shifted))["scene_features"], r_mean["scene_features"])
    ref = oracle.evaluate_scene(sc, small_state, k=16, n_iter=3, min_visible=50)
    assert set(r_mean) == {"scene_features", "text_features", "logit_scale"}
    assert r_mean["scene_features"].shape == (n, 128) and r_cons["scene_features"].shape == (n, 128)
    assert (r_mean["scene_features"].cpu() - ref["scene_features"]).abs().max().item() < 1e-3
    # consensus lifting gives different (but finite, non-degenerate) features on the same surface
    assert torch.isfinite(r_cons["scene_features"]).all() and r_cons["scene_features"].abs().sum() > 0


def test_ape_variant_d256_tuple_api_matches_oracle():
    """`lift_ape_features` (models/affinity_module.py:736-993; selected at :1497): the consensus fusion on 256-d
    instance embeddings, then the whole purification at D = 256 (student 262 -> ..., tcgen05 pooling in its one-pass
    256-column form) through the reference's tuple surface, against the oracle's restatement of the same lines."""
    from geopurify_b200.affinity_module import SonataXAffinityTrainer
    from geopurify_b200.student import AffinityPredictor, load_numpy_state
    from oracle import oracle
    sc = synth.make_scene(n_points=9000, n_views=5, image_dim=(80, 60), seed=13, room=(2.0, 1.6, 2.0), n_boxes=2,
                          cut=2, q_range=(4, 12), feat_dim=256, n_classes=7)
    state = synth.make_student_state(input_dim=262, hidden_dim=128, embed_dim=128, seed=13)
    W, H = sc.image_dim
    kmap = oracle.mapper_intrinsics(sc.intrinsics, sc.image_dim)
    maps = np.stack([oracle.project_view(sc.w2c[v].T.copy(), sc.coords.astype(np.float64), sc.depth[v], kmap, (W, H),
                                         sc.vis_thres, sc.cut)[0] for v in range(sc.n_views)])
    kept, rows, cols, pids = oracle.accept_views(maps, 50)
    n = sc.n_points
    batch = [None] * 20
    batch[0] = torch.as_tensor(sc.coords)
    batch[1] = torch.as_tensor(sc.voxel_coords).float()
    batch[2] = torch.as_tensor(sc.inds_reconstruct)
    batch[3] = torch.as_tensor(sc.labels)
    batch[12] = torch.cat([torch.as_tensor(r) for r in rows])
    batch[13] = torch.cat([torch.as_tensor(c) for c in cols])
    batch[14] = torch.cat([torch.stack([torch.full((n,), i), torch.as_tensor(maps[v][:, 2])], 1)
                           for i, v in enumerate(kept)])
    batch[19] = torch.as_tensor(sc.gauss)
    idm = torch.as_tensor(sc.idmaps[kept]).cuda()
    tab = torch.as_tensor(sc.tables[kept]).cuda()
    model = SonataXAffinityTrainer(mask_features=lambda b: (idm, tab), lifting="ape", text_features=sc.text,
                                   logit_scale=sc.logit_scale, feature_dim=256, K=16, n_pool=3,
                                   student=load_numpy_state(AffinityPredictor(262, 128, 128), state))
    assert model.use_ape and not model.use_lseg
    got = model.evaluate_scene(tuple(batch), keep_intermediates=True)
    assert got["scene_features"].shape == (n, 256)

    vpf = [sc.tables[v][sc.idmaps[v][rows[i], cols[i]].astype(np.int64)] for i, v in enumerate(kept)]
    lifted, seen = oracle.lift_consensus(n, sc.coords, vpf, pids, sc.text, sc.logit_scale, 256)
    assert (model.last_lifted.cpu() - lifted).abs().max().item() < 1e-5
    vin = torch.cat([oracle.scatter_mean(lifted, sc.inds_reconstruct),
                     oracle.scatter_mean(sc.gauss[:, :6], sc.inds_reconstruct)], 1)
    emb = oracle.student_forward(vin, sc.voxel_coords, state)
    nbr = oracle.knn_c(sc.voxel_coords, 16)
    w = oracle.affinity_weights(emb, nbr, 20.0)
    want = oracle.pool(vin, nbr, w, 3)[torch.as_tensor(sc.inds_reconstruct)][:, :256]
    err = (got["scene_features"].cpu() - want).abs().max().item()
    print(f"APE variant D=256: scene features max-abs {err:.2e}")
    assert err < 1e-3
    with pytest.raises(ValueError):
        SonataXAffinityTrainer(text_features=sc.text, lifting="nope")
