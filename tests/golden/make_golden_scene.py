"""Golden vectors for the whole inference path from the REFERENCE'S OWN CODE.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_scene.py

What executes unmodified from /root/reference:
  * `models/utils/fusion_util.py:PointCloudToImageMapper.compute_mapping`   (per view, as the loader calls it)
  * `dataset/voxelizer.py:Voxelizer.voxelize`                                (scene voxels + inds_reconstruct)
  * `dataset/data_loader_ablation.py:scene_based_collate_fn`                 (the 20-tuple `batch_data`)
  * `models/affinity_module.py:SonataXAffinityTrainer.evaluate_scene` -> `lift_lseg_features` (gather, mean over
    views, KDTree fill), scatter-mean, `AffinityPredictor` (the reference's class: layer wiring, BN/ReLU/residual
    order), normalise, kNN K=96 with the self column dropped, softmax(20 cos), `torch.sparse.mm` x19, point gather.
What is stood in for (tests/golden/ref_stubs, written from the libraries' documented semantics): MinkowskiEngine
(SparseTensor / 3^3 convolution / BatchNorm / ReLU), faiss.IndexFlatL2, torch_scatter.scatter_mean; the LSeg network
is replaced by a table of seeded per-view feature maps (the 2-D VLM is outside the path).  sklearn's KDTree is real.
The per-view item is assembled as `data_loader_ablation.py:242-392` does (those lines sit inside a file-reading
method and cannot be called on their own).

Only inputs (or the seeds that make them) and outputs are stored.
"""
import collections
import collections.abc
import os
import sys

import numpy as np
import torch

collections.Sequence = collections.abc.Sequence
collections.Iterable = collections.abc.Iterable
sys.dont_write_bytecode = True
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, "/root/reference")
import ref_stubs  # noqa: E402

ref_stubs.install()
from dataset.data_loader_ablation import scene_based_collate_fn  # noqa: E402
from dataset.voxelizer import Voxelizer  # noqa: E402
from models.affinity_module import AffinityPredictor, SonataXAffinityTrainer  # noqa: E402
from models.utils.fusion_util import PointCloudToImageMapper  # noqa: E402
from geopurify_b200 import synth  # noqa: E402

SCENE = dict(n_points=2500, n_views=6, image_dim=(80, 60), seed=23, room=(2.2, 1.8, 2.2), n_boxes=2, cut=3,
             q_range=(4, 9), feat_dim=512, n_classes=20)
STUDENT_SEED = 5557
ROWS = 384     # rows of the [N,512] outputs stored in full; every row also gets a float64 checksum


class TableLSeg:
    """Stands where `LSeg_MultiEvalModule` stands: returns the next B views' [B,512,H,W] maps."""

    def __init__(self, maps):
        self.maps, self.cursor = maps, 0

    def eval(self):
        return self

    def forward(self, x, label_set=""):
        b = x.shape[0]
        out = self.maps[self.cursor:self.cursor + b]
        self.cursor += b
        return out


def view_item(v, sc, mapper, voxelizer, scene_vox, scene_rec):
    """One dataset item, following dataset/data_loader_ablation.py:242-392."""
    H, W = sc.image_dim[1], sc.image_dim[0]
    locs_in = sc.coords.astype(np.float32)
    labels_in = sc.labels.astype(np.int64)
    point_features = sc.gauss.astype(np.float32)
    mapping = np.ones([locs_in.shape[0], 4], dtype=int)
    mapping[:, 1:4], _ = mapper.compute_mapping(sc.w2c[v].T.copy(), locs_in, sc.depth[v])       # :248-253
    if mapping[:, 3].sum() == 0:                                                                  # :254
        return None
    mask = mapping[:, 3]
    label_3d = labels_in[mask == 1].copy()
    feature_3d = point_features[mask == 1].copy()
    locals_3d = locs_in[mask == 1].copy()
    unique_map = mapping.copy()
    locs, feats, _, inds_reconstruct = voxelizer.voxelize(locals_3d, feature_3d, label_3d)       # :324
    coords = torch.from_numpy(locs).int()
    coords = torch.cat((torch.ones(coords.shape[0], 1, dtype=torch.int), coords), dim=1)
    feats = torch.ones(coords.shape[0], 3)
    x_label = torch.from_numpy(mapping[:, 1][mapping[:, 1] != 0]).long()                          # :348-351
    y_label = torch.from_numpy(mapping[:, 2][mapping[:, 2] != 0]).long()
    locals_3d_t = torch.from_numpy(locals_3d).float()
    locals_3d_t = torch.cat((torch.ones(locals_3d_t.shape[0], 1, dtype=torch.float), locals_3d_t), dim=1)
    img = torch.zeros(H, W, 3)
    return (torch.from_numpy(locs_in).float(), torch.from_numpy(scene_vox).float(),
            torch.from_numpy(scene_rec).long(), torch.from_numpy(labels_in).long(), locals_3d_t, coords, feats,
            torch.from_numpy(feature_3d).float(), torch.from_numpy(label_3d).long(), torch.zeros(label_3d.shape[0]),
            torch.zeros(H, W, dtype=torch.long), img, x_label, y_label, torch.from_numpy(mask).bool(),
            torch.from_numpy(inds_reconstruct).long(), torch.from_numpy(unique_map).long(),
            torch.from_numpy(mapping), None, torch.from_numpy(point_features).float())


class TableXDecoder:
    """Stands where `self.xdecoder_teacher.model` stands: `forward_seg_all` returns the next view's head outputs
    (models/affinity_module.py:519 reads pred_logits / pred_masks / mask_embed / text_embed / logit_scale)."""

    def __init__(self, masks, logits, embed, text, logit_scale):
        self.masks, self.logits, self.embed = masks, logits, embed
        self.text, self.logit_scale, self.cursor = text, logit_scale, 0
        self.model = self

    def forward_seg_all(self, inputs):
        v = self.cursor
        self.cursor += 1
        return None, {"pred_logits": self.logits[v:v + 1], "pred_masks": self.masks[v:v + 1],
                      "mask_embed": self.embed[v:v + 1], "text_embed": self.text, "logit_scale": self.logit_scale}


def checksum(x):
    x = x.double()
    return torch.stack([x.sum(1), (x * x).sum(1)], 1).numpy()


def main():
    sc = synth.make_scene(**SCENE)
    W, H = sc.image_dim
    mapper = PointCloudToImageMapper([W, H], sc.vis_thres, sc.cut, sc.intrinsics)
    voxelizer = Voxelizer(voxel_size=0.02, clip_bound=None, use_augmentation=False)
    locs_in = sc.coords.astype(np.float32)
    scene_vox, _, _, scene_rec = voxelizer.voxelize(locs_in, sc.gauss.astype(np.float32), sc.labels.astype(np.int64))
    items = [view_item(v, sc, mapper, voxelizer, scene_vox, scene_rec) for v in range(sc.n_views)]
    kept = [v for v, it in enumerate(items) if it is not None]
    batch_data = scene_based_collate_fn(items)

    maps = torch.stack([torch.from_numpy(sc.dense_map(v)).permute(2, 0, 1) for v in kept])     # [V,512,H,W]
    t = SonataXAffinityTrainer.__new__(SonataXAffinityTrainer)
    torch.nn.Module.__init__(t)
    t.device, t.use_lseg, t.use_ape = "cpu", True, False
    t.lseg_evaluator, t.lseg_normalize = TableLSeg(maps), (lambda x: x)
    t.text_features, t.logit_scale = torch.from_numpy(sc.text).float(), torch.tensor(float(sc.logit_scale))
    student = AffinityPredictor(input_dim=518, embed_dim=128, hidden_dim=512)
    state = synth.make_student_state(518, 512, 128, seed=STUDENT_SEED)
    sd = student.state_dict()
    for k in sd:
        if k in state:
            sd[k] = torch.as_tensor(np.asarray(state[k]))
    missing = [k for k in sd if k not in state and not k.endswith("num_batches_tracked")]
    assert not missing, missing
    student.load_state_dict(sd)
    t.affinity_student = student.eval()

    with torch.no_grad():
        lifted, _, _ = t.lift_lseg_features(batch_data)
        t.lseg_evaluator.cursor = 0
        out = t.evaluate_scene(batch_data)
    feats = out["scene_features"]
    rows = np.sort(np.random.default_rng(1).choice(feats.shape[0], ROWS, replace=False))
    logits = torch.nn.functional.normalize(feats, dim=-1) @ torch.nn.functional.normalize(out["text_features"], dim=-1).t()
    top2 = logits.topk(2, dim=1).values
    np.savez_compressed(
        os.path.join(HERE, "scene_lseg.npz"),
        scene_kwargs=np.array(repr(SCENE)), student_seed=STUDENT_SEED, kept=np.array(kept),
        coords=locs_in, coords_checksum=np.float64(sc.coords.astype(np.float64).sum()),
        voxel_coords=scene_vox.astype(np.int32), inds_reconstruct=scene_rec.astype(np.int64),
        x_labels=batch_data[12].numpy().astype(np.int32), y_labels=batch_data[13].numpy().astype(np.int32),
        mask_2ds=batch_data[14].numpy().astype(np.uint8),
        rows=rows, lifted_rows=lifted[rows].numpy(), lifted_checksum=checksum(lifted),
        scene_rows=feats[rows].numpy(), scene_checksum=checksum(feats),
        labels=logits.argmax(1).numpy().astype(np.int32), margin=(top2[:, 0] - top2[:, 1]).numpy(),
        stubbed=np.array(sorted(set(s.split(".")[0] for s in ref_stubs.stubbed))))
    validate_case(sc, t, batch_data, out)
    xdecoder_case(sc, t, batch_data, kept)
    mining_case(t)
    print("kept views", kept, "points", feats.shape[0], "voxels", scene_vox.shape[0],
          "stubbed:", sorted(set(s.split(".")[0] for s in ref_stubs.stubbed)))
    print("feature range", float(feats.min()), float(feats.max()), "min margin", float((top2[:, 0] - top2[:, 1]).min()))


ZERO_SEED, ZERO_FRAC = 3, 0.1
BASE, NOVEL = list(range(0, 12)), list(range(12, 20))


def validate_case(sc, trainer, batch_data, first_result):
    """A12: `run/validation.py:validate(val_loader, model)` executed unmodified over a two-scene loader.  Scene 0 is
    the scene above (the model's `evaluate_scene` result is the reference trainer's); scene 1 is the same scene with
    10 % of the feature rows zeroed, which sends those points through the KDTree fill of :421-433 (it searches on
    `scene_coords[:, 1:4]`, i.e. the y and z columns only).  Module globals the function reads (`args`, `logger`) are
    set from outside; `intersectionAndUnionGPU` is wrapped to record its arguments and results, and `Tensor.cuda` is
    the identity while it runs (no GPU here)."""
    import logging
    import tempfile
    import types
    import run.validation as rv
    n = first_result["scene_features"].shape[0]
    zero = np.sort(np.random.default_rng(ZERO_SEED).choice(n, int(ZERO_FRAC * n), replace=False))
    second = dict(first_result)
    second["scene_features"] = first_result["scene_features"].clone()
    second["scene_features"][zero] = 0.0
    results = [first_result, second]

    class Replay:
        calls = 0

        def eval(self):
            return self

        def evaluate_scene(self, batch_data, vis_prefix=None):
            r = results[self.calls]
            Replay.calls += 1
            return {k: (v.clone() if torch.is_tensor(v) else v) for k, v in r.items()}

    class Loader(list):
        dataset = types.SimpleNamespace(data_paths=["scene0000_00_vh_clean_2.pth", "scene0000_01_vh_clean_2.pth"])

    recorded = []
    ref_metric = rv.intersectionAndUnionGPU

    def recorder(output, target, k, ignore):
        i, u, t = ref_metric(output, target, k, ignore)
        recorded.append((output.clone().numpy(), i.numpy(), u.numpy(), t.numpy()))
        return i, u, t

    rv.args = types.SimpleNamespace(test_classes=20, test_ignore_label=[sc.ignore_label], multiprocessing_distributed=False,
                                    category_split=types.SimpleNamespace(base_category=BASE, novel_category=NOVEL))
    rv.logger = logging.getLogger("reference-validate")
    rv.intersectionAndUnionGPU = recorder
    cuda, cwd = torch.Tensor.cuda, os.getcwd()
    torch.Tensor.cuda = lambda self, *a, **kw: self
    try:
        with tempfile.TemporaryDirectory() as tmp:
            os.chdir(tmp)                                   # validate() creates its visualisation directory in the cwd
            miou_base, miou_novel = rv.validate(Loader([batch_data, batch_data]), Replay())
    finally:
        os.chdir(cwd)
        torch.Tensor.cuda = cuda
        rv.intersectionAndUnionGPU = ref_metric
    assert len(recorded) == 2
    np.savez_compressed(
        os.path.join(HERE, "validate_two_scenes.npz"),
        zero_rows=zero, base=np.array(BASE), novel=np.array(NOVEL), ignore=np.array([sc.ignore_label]),
        pred=np.stack([r[0] for r in recorded]).astype(np.int32), inter=np.stack([r[1] for r in recorded]),
        union=np.stack([r[2] for r in recorded]), target=np.stack([r[3] for r in recorded]),
        miou_base=np.float64(miou_base), miou_novel=np.float64(miou_novel))
    print("validate(): mIoU base/novel", miou_base, miou_novel, "filled predictions", len(zero))


XDEC = dict(n_queries=14, seed=7)


def xdecoder_case(sc, t, batch_data, kept):
    """A5: `lift_xdecoder_features` (models/affinity_module.py:455-714) with the X-Decoder forward replaced by seeded
    head outputs.  `mappings` is handed over per visible pair: the function indexes it with the per-pair view mask
    (:504), which the [V*N,4] tensor of `scene_based_collate_fn` would not survive; it is not used afterwards."""
    import types
    W, H = sc.image_dim
    masks, logits, embed = synth.make_mask_predictions(len(kept), XDEC["n_queries"], (H, W), sc.text, 512, XDEC["seed"])
    t.use_lseg = False
    t.cfg = types.SimpleNamespace(all_label=list(range(sc.text.shape[0])), mask_shape=(H, W), use_ape=False)
    t.xdecoder_teacher = TableXDecoder(torch.from_numpy(masks), torch.from_numpy(logits), torch.from_numpy(embed),
                                       torch.from_numpy(sc.text).float(), torch.tensor(float(sc.logit_scale)))
    bd = list(batch_data)
    bd[17] = bd[4].clone()
    with torch.no_grad():
        lifted, text_n, scale = t.lift_xdecoder_features(tuple(bd))
    rows = np.sort(np.random.default_rng(2).choice(lifted.shape[0], ROWS, replace=False))
    zero_rows = int((lifted.abs().sum(1) == 0).sum())
    np.savez_compressed(
        os.path.join(HERE, "scene_xdecoder.npz"),
        scene_kwargs=np.array(repr(SCENE)), kept=np.array(kept), n_queries=XDEC["n_queries"], mask_seed=XDEC["seed"],
        coords=sc.coords.astype(np.float32),
        x_labels=batch_data[12].numpy().astype(np.int32), y_labels=batch_data[13].numpy().astype(np.int32),
        mask_2ds=batch_data[14].numpy().astype(np.uint8),
        rows=rows, lifted_rows=lifted[rows].numpy(), lifted_checksum=checksum(lifted),
        text_normalised_checksum=checksum(text_n))
    print("xdecoder lifting:", tuple(lifted.shape), "all-zero rows", zero_rows, "logit_scale", float(scale))


MINE = dict(n=1500, d=64, seed=11, torch_seed=5557)


def mining_inputs():
    """Seeded inputs of the mining case (tests rebuild them the same way): points, clustered features, their kNN."""
    from oracle import oracle
    xyz, feat = synth.make_mining_inputs(MINE["n"], MINE["d"], MINE["seed"])
    return xyz, feat, oracle.knn_points(xyz, 96)


def mining_case(t):
    """F4: the reference's `sample_contrastive_pairs_hybrid` (models/affinity_module.py:1099-1136) executed unmodified;
    its `torch.randperm` anchors are stored with the outputs."""
    xyz, feat, nbr = mining_inputs()
    t.num_anchors_per_scene, t.num_negatives_per_anchor = 4096, 63
    torch.manual_seed(MINE["torch_seed"])
    with torch.no_grad():
        anchor, pos, neg = t.sample_contrastive_pairs_hybrid(torch.from_numpy(feat), torch.from_numpy(nbr))
    np.savez_compressed(os.path.join(HERE, "mining_hybrid.npz"), params=np.array(repr(MINE)),
                        feat_checksum=np.float64(feat.astype(np.float64).sum()), anchors=anchor.numpy().astype(np.int32),
                        positive=pos.numpy().astype(np.int32), negatives=neg.numpy().astype(np.int32))
    print("mining:", tuple(anchor.shape), tuple(pos.shape), tuple(neg.shape))


if __name__ == "__main__":
    main()
