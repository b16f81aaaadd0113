"""Full-size checks (BASELINE.json configs[1] shape: N=150k, V=50, 240x320, D=512, K=96, T=19, hidden 512).

Two kinds: (i) the CUDA path against the CPU oracle at the headline size -- visibility / pixels / view acceptance /
kNN / labels / histograms bit for bit, embeddings and pooled features within tolerance, mIoU identical, for bf16 AND
fp32 dense 2-D maps, at C = 20 and C = 200 (the oracle needs about a minute of host time: OpenMP brute-force kNN,
per-offset matmul student, spmm_ref.c pooling); (ii) properties that need no oracle: exact kNN against a brute force
on sampled queries, (d^2, index) ordering, pooling as a row-stochastic operator, histogram conservation, determinism;
(iii) the multi-slab branch of the tcgen05 student (M > 131072 voxels) against the fp32 SIMT student."""
import numpy as np
import pytest
import torch

from geopurify_b200 import synth

pytestmark = pytest.mark.gpu


def _bf16_round(a):
    return torch.as_tensor(a).to(torch.bfloat16).float().numpy()


@pytest.fixture(scope="module")
def full_scene():
    sc = synth.make_scene(n_points=150_000, n_views=50, image_dim=(320, 240), feat_dim=512, n_classes=20, seed=5557)
    # 2-D features exactly representable in bf16, so that the bf16-map run, the fp32-map run and the oracle all
    # read the same numbers (the bench's `feat_dtype_2d: bf16` rounds the same way on upload)
    sc.tables = _bf16_round(sc.tables)
    return sc


@pytest.fixture(scope="module")
def full_run(full_scene):
    from geopurify_b200 import _lib
    from geopurify_b200.affinity_module import SonataXAffinityTrainer, scene_inputs_from_synth
    from geopurify_b200.student import AffinityPredictor, load_numpy_state
    _lib.require_cuda()
    sc = full_scene
    state = synth.make_student_state()
    student = load_numpy_state(AffinityPredictor(518, 128, 512), state)
    model = SonataXAffinityTrainer(text_features=sc.text, logit_scale=100.0, student=student, feature_dim=512)
    s = scene_inputs_from_synth(sc, feat_dtype=torch.bfloat16)
    ign = torch.tensor([sc.ignore_label], dtype=torch.int64, device="cuda")
    out = model.run_scene(s, ignore=ign, keep_intermediates=True)
    torch.cuda.synchronize()
    return dict(sc=sc, model=model, s=s, out=out, ign=ign, state=state, student=student)


@pytest.fixture(scope="module")
def full_oracle(full_scene):
    """oracle.evaluate_scene at the headline size (K=96, T=19, hidden 512) -- about a minute on the host cores."""
    import os
    import time
    from oracle import oracle
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    torch.set_num_threads(cores)
    stages = {}
    t0 = time.perf_counter()
    ref = oracle.evaluate_scene(full_scene, synth.make_student_state(), k=96, n_iter=19, stages=stages)
    print(f"full-size oracle: {time.perf_counter() - t0:.1f} s on {cores} threads "
          + " ".join(f"{k}={v:.1f}" for k, v in stages.items()))
    return ref


def _compare_with_oracle(tag, sc, model, out, ref, text, n_classes):
    """Everything the north star lists, hard: integer outputs ==, float outputs within bound, mIoU identical."""
    from geopurify_b200.validation import mean_iou
    from oracle import oracle
    order = model.last["index"].order.long().cpu()
    assert (model.last_seen.cpu().numpy().astype(bool) == ref["seen"].numpy()).all(), "seen flags differ"
    keep = model.last_keep.cpu().numpy().astype(bool)
    assert sorted(np.nonzero(keep)[0].tolist()) == list(ref["kept"]), "accepted views differ"
    assert (model.last["knn_orig"].cpu().numpy().astype(np.int64) == ref["knn"]).all(), "kNN indices differ"
    e_lift = (model.last_lifted.cpu() - ref["lifted"]).abs().max().item()
    e_vin = (model.last["voxel_in"][:, :518].cpu() - ref["voxel_in"][order]).abs().max().item()
    e_emb = (model.last["embed"].cpu() - ref["embed"][order]).abs().max().item()
    e_w = (model.last["weights"].cpu() - ref["weights"][order]).abs().max().item()
    e_pool = (model.last["pooled"].cpu() - ref["pooled_voxels"][order][:, :512]).abs().max().item()
    assert e_lift == 0.0, f"{tag}: lifted features differ by {e_lift} (A4 is bit-exact on identical inputs)"
    assert e_vin <= 1e-6, f"{tag}: voxel means differ by {e_vin}"
    assert e_emb < 5e-5, f"{tag}: embedding max-abs {e_emb}"
    assert e_w < 5e-4, f"{tag}: softmax weights max-abs {e_w}"
    assert e_pool < 1e-3, f"{tag}: pooled max-abs {e_pool} (north star bound 1e-3)"
    # classification against the oracle's pooled features with this text matrix
    ref_pred, ref_logits = oracle.classify(ref["scene_features"], text, 100.0, sc.coords[:, 1:4])
    i_, u_, t_ = oracle.intersection_and_union(ref_pred, torch.as_tensor(sc.labels), n_classes, (sc.ignore_label,))
    ref_pred = ref_pred.clone()
    ref_pred[torch.as_tensor(sc.labels) == sc.ignore_label] = sc.ignore_label
    pred = out["pred"].cpu()
    bad = torch.nonzero(pred != ref_pred).flatten()
    top2 = ref_logits.double().topk(2, dim=1).values
    margin = (top2[:, 0] - top2[:, 1])
    print(f"{tag}: M={order.numel()} lifted {e_lift:.1e} vin {e_vin:.1e} emb {e_emb:.2e} weights {e_w:.2e} pooled {e_pool:.2e} "
          f"label mismatches {bad.numel()} of {pred.numel()}"
          + (f" (oracle top-2 logit margins there: {margin[bad][:8].tolist()})" if bad.numel() else ""))
    assert bad.numel() == 0, f"{tag}: {bad.numel()} labels differ from the oracle"
    h = out["hist"].cpu().numpy()
    assert (h[0] == i_.numpy()).all() and (h[1] == u_.numpy()).all() and (h[2] == t_.numpy()).all(), "histograms"
    assert mean_iou(h[0], h[1], h[2]) == oracle.miou(i_.numpy(), u_.numpy(), t_.numpy()), "mIoU not identical"


def test_fullsize_projection_matches_oracle(full_run, full_oracle):
    """A1 + A3 at N=150k x 50 views: every (row, col, visible) triple equals the oracle's."""
    from geopurify_b200 import ops
    s = full_run["s"]
    W, H = s.image_dim
    pix, counts = ops.project_views(s.coords, s.w2c, s.intr, s.depth, H, W, s.cut, s.vis_thres)
    got = ops.pix_to_mapping(pix, W).cpu().numpy()
    assert got.shape == full_oracle["mappings"].shape
    assert (got == full_oracle["mappings"]).all(), "projected pixels / visibility differ"
    assert (counts.cpu().numpy() == full_oracle["mappings"][:, :, 2].sum(1)).all()


def test_fullsize_bf16_maps_match_oracle(full_run, full_oracle):
    r = full_run
    _compare_with_oracle("bf16 maps, C=20", r["sc"], r["model"], r["out"], full_oracle, r["sc"].text, 20)


def test_fullsize_fp32_maps_match_oracle(full_run, full_oracle):
    from geopurify_b200.affinity_module import SonataXAffinityTrainer, scene_inputs_from_synth
    r = full_run
    sc = r["sc"]
    model = SonataXAffinityTrainer(text_features=sc.text, logit_scale=100.0, student=r["student"], feature_dim=512)
    s = scene_inputs_from_synth(sc, feat_dtype=torch.float32)
    out = model.run_scene(s, ignore=r["ign"], keep_intermediates=True)
    torch.cuda.synchronize()
    _compare_with_oracle("fp32 maps, C=20", sc, model, out, full_oracle, sc.text, 20)
    assert torch.equal(out["labels_voxel"], r["out"]["labels_voxel"]), "bf16 and fp32 maps of equal values disagree"
    assert torch.equal(out["pooled"], r["out"]["pooled"])


def test_fullsize_200_classes_match_oracle(full_run, full_oracle):
    """BASELINE.json configs[2]: same geometry, 200-class text embeddings (the tcgen05 logits kernel's 256-column
    instance + its fp32 fix-up of near ties)."""
    from geopurify_b200.affinity_module import SonataXAffinityTrainer
    r = full_run
    sc = r["sc"]
    rng = np.random.default_rng(5557 + 200)
    text = rng.standard_normal((200, 512)).astype(np.float32)
    text /= np.linalg.norm(text, axis=1, keepdims=True)
    model = SonataXAffinityTrainer(text_features=text, logit_scale=100.0, student=r["student"], feature_dim=512)
    out = model.run_scene(r["s"], ignore=r["ign"], keep_intermediates=True)
    torch.cuda.synchronize()
    _compare_with_oracle("bf16 maps, C=200", sc, model, out, full_oracle, text, 200)


def test_student_multi_slab_matches_fp32_simt():
    """M > 131072 voxels takes the tcgen05 student's multi-slab branch (slab-relative pair / tile tables, staging
    reuse, neighbours across the slab boundary).  The bench's 171k-point scenes and the Matterport shape run it;
    here it is compared with the fp32 SIMT student on the same rows, and labels downstream must agree."""
    from geopurify_b200 import _lib, ops
    from geopurify_b200.student import AffinityPredictor, load_numpy_state
    _lib.require_cuda()
    sc = synth.make_scene(n_points=185_000, n_views=4, image_dim=(80, 60), feat_dim=512, n_classes=20, seed=77,
                          cut=2, q_range=(4, 12))
    m = sc.voxel_coords.shape[0]
    assert m > 131072 + 4096, f"scene too small for two slabs: M={m}"
    vox = torch.as_tensor(sc.voxel_coords.astype(np.int32)).cuda()
    index = ops.VoxelIndex(vox)
    nbr27 = index.kernel_map()
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.zeros((m, 544), dtype=torch.float32, device="cuda")
    x[:, :518] = torch.randn((m, 518), device="cuda", generator=g) * 0.05
    student = load_numpy_state(AffinityPredictor(518, 128, 512), synth.make_student_state())
    student.precision = "bf16x3"
    tc = student.forward_sorted(x, nbr27, normalize=True)
    tc2 = student.forward_sorted(x, nbr27, normalize=True)
    student.precision = "fp32"
    ref = student.forward_sorted(x, nbr27, normalize=True)
    torch.cuda.synchronize()
    assert torch.equal(tc, tc2), "tcgen05 student is not run-to-run deterministic"
    err = (tc - ref).abs().max().item()
    # rows on both sides of the slab boundary, and rows whose 3^3 neighbourhood crosses it
    cross = ((nbr27 >= 131072).any(1) & (torch.arange(m, device="cuda") < 131072)) | \
            (((nbr27 >= 0) & (nbr27 < 131072)).any(1) & (torch.arange(m, device="cuda") >= 131072))
    print(f"multi-slab student: M={m} max-abs {err:.2e}, {int(cross.sum())} rows with neighbours across the boundary, "
          f"max-abs there {(tc - ref)[cross].abs().max().item() if bool(cross.any()) else 0.0:.2e}")
    assert err < 2e-5, f"tcgen05 (2 slabs) vs fp32 SIMT student: {err}"


def test_fullsize_knn_exact_on_sampled_queries(full_run):
    model, s = full_run["model"], full_run["s"]
    knn = model.last["knn_orig"].long()                     # [M,96] original numbering
    vox = s.voxel_coords.long()
    m, K = knn.shape
    assert K == 96 and int(knn.min()) >= 0 and int(knn.max()) < m
    assert not bool((knn == torch.arange(m, device="cuda")[:, None]).any()), "self in a neighbour list"
    d2 = ((vox[knn] - vox[:, None, :]) ** 2).sum(-1)
    key = d2 * m + knn
    assert bool((key[:, 1:] > key[:, :-1]).all()), "lists not strictly ascending in (d^2, index)"
    g = torch.Generator(device="cuda").manual_seed(0)
    q = torch.randint(0, m, (256,), device="cuda", generator=g)
    dq = ((vox[q][:, None, :] - vox[None, :, :]) ** 2).sum(-1)          # [256, M] exact integers
    kq = dq * m + torch.arange(m, device="cuda")[None, :]
    kq[torch.arange(256, device="cuda"), q] = torch.iinfo(torch.int64).max   # drop self
    want = torch.sort(kq, dim=1).values[:, :K] % m
    assert torch.equal(want, knn[q]), "kNN differs from the brute force"


def test_fullsize_pooling_is_row_stochastic(full_run):
    from geopurify_b200 import ops
    model = full_run["model"]
    w, nbr, vin, pooled = (model.last[k] for k in ("weights", "knn_sorted", "voxel_in", "pooled"))
    assert torch.allclose(w.sum(1), torch.ones(w.shape[0], device="cuda"), atol=2e-6) and bool((w >= 0).all())
    ones = torch.ones((w.shape[0], 512), device="cuda")
    assert (ops.pool(ones, 512, nbr, w, 19) - 1.0).abs().max().item() < 2e-5      # A^19 1 = 1
    lo, hi = vin[:, :512].min(0).values, vin[:, :512].max(0).values              # convex combinations stay in range
    assert bool((pooled >= lo - 1e-5).all()) and bool((pooled <= hi + 1e-5).all())
    tc = ops.pool(vin, 512, nbr, w, 19, tensor_core=True)
    rw = ops.pool(vin, 512, nbr, w, 19, tensor_core=False)
    assert (tc - rw).abs().max().item() < 1e-5, "tcgen05 pooling vs fp32 row-wise pooling"


def test_fullsize_histograms_conserve_points_and_run_is_deterministic(full_run):
    sc, model, s, out, ign = (full_run[k] for k in ("sc", "model", "s", "out", "ign"))
    h = out["hist"].cpu().numpy()
    valid = int((sc.labels != sc.ignore_label).sum())
    assert h[2].sum() == valid                                   # target histogram counts every labelled point
    assert (h[0] <= h[2]).all() and (h[0] <= h[1]).all()         # intersection <= target, union
    assert h[1].sum() == 2 * valid - h[0].sum()                  # |out| + |tgt| - |inter| with |out| = |tgt| = valid
    again = model.run_scene(s, ignore=ign, keep_intermediates=False)
    assert torch.equal(again["labels_voxel"], out["labels_voxel"]) and torch.equal(again["hist"], out["hist"])
    assert torch.equal(again["pooled"], out["pooled"]), "run-to-run bit differences"


def test_fullsize_lifting_and_fill(full_run):
    model = full_run["model"]
    lifted, seen = model.last_lifted, model.last_seen
    assert 0.3 < float(seen.float().mean()) < 1.0
    assert bool((lifted.abs().sum(1) > 0).all()), "a point kept a zero feature after the fill"
    assert bool(torch.isfinite(lifted).all())


def test_fullsize_workspace_guard_bands(full_run):
    """Headline-size run (single 131072-voxel student slab, 1018 pooling tiles, 150k-point fill grid) with canaries
    around every workspace -- the substitute for compute-sanitizer memcheck, which is closed on the build pool."""
    from geopurify_b200 import ops
    r = full_run
    was = ops._GUARD
    ops.guard_workspaces(True)
    try:
        again = r["model"].run_scene(r["s"], ignore=r["ign"])
        assert ops.check_guards() >= 6
        assert torch.equal(again["labels_voxel"], r["out"]["labels_voxel"])
    finally:
        ops.guard_workspaces(was)
