"""Host-side mirror of the reference's `SonataXAffinityTrainer` inference surface on the CUDA path.

Reference: models/affinity_module.py -- `lift_lseg_features` (:348-453), `evaluate_scene` (:1490-1608).
The 2-D VLM (X-Decoder / LSeg) and the Sonata teacher are out of scope: the
per-view 2-D features are supplied by a `view_features` provider as channels-last device tensors.

Two entry styles:
  * `evaluate_scene(batch_data)`        -- the reference's positional 20-tuple (CPU tensors), returns the same
                                           dict {"scene_features", "text_features", "logit_scale"};
  * `run_scene(SceneInputs)`            -- the B200-first form: cameras + depth instead of precomputed pixel
                                           lists, fused projection/gather, everything stays on the device,
                                           returns labels + IoU histograms.
Internally voxels are processed in Hilbert-cell order; every exposed index tensor is in the reference's order.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from . import _lib as L
from . import ops
from .student import AffinityPredictor


@dataclass
class SceneInputs:
    """Device-resident inputs of one scene for `run_scene` (data layout)."""
    coords: torch.Tensor           # [N,3] float64 or float32 world xyz (float64 is what the .pth files hold)
    coords_f32: torch.Tensor       # [N,3] float32 (scene_coords of the batch tuple; hole fills use it)
    gauss: torch.Tensor            # [N,6] float32 rgb | normal
    voxel_coords: torch.Tensor     # [M,3] int32
    p2v: torch.Tensor              # [N] int32 point -> voxel (scene_inds_reconstruct)
    extent: tuple                  # (ex, ey, ez) exclusive bound of voxel_coords (host ints)
    w2c: torch.Tensor              # [V,16] float64
    intr: torch.Tensor             # [V,4] float64
    depth: Optional[torch.Tensor]  # [V,H,W] float64/float32
    feat: torch.Tensor             # [V,H,W,D] bf16/float32 channels-last 2-D features
    image_dim: tuple               # (W, H)
    cut: int
    vis_thres: float
    labels: Optional[torch.Tensor] = None   # [N] int64
    idmaps: Optional[torch.Tensor] = None   # [V,H,W] uint8  compact 2-D form (consensus lifting): mask id per pixel
    tables: Optional[torch.Tensor] = None   # [V,Q,D] float32 mask embeddings, all-zero row = no mask


class SonataXAffinityTrainer:
    """Inference half of the reference class of the same name.  (`GeoPurifyB200` is an alias.)"""

    K = 96                        # models/affinity_module.py:1492
    affinity_sharpen_factor = 20  # :1493
    n_pool = 19                   # :1584-1587

    def __init__(self, cfg=None, text_features=None, logit_scale=100.0, view_features=None, student=None,
                 feature_dim=512, K=None, n_pool=None, sharpen=None, min_visible=400, fill_cell=0.05,
                 device="cuda", mask_features=None, lifting=None):
        L.require_cuda()
        self.cfg = cfg
        self.device = torch.device(device)
        self.feature_dim = feature_dim
        self.affinity_student = student if student is not None else AffinityPredictor(
            input_dim=feature_dim + 6, embed_dim=128, hidden_dim=512)          # :266-272
        self.text_features = None if text_features is None else torch.as_tensor(text_features).float().to(self.device)
        self.logit_scale = float(logit_scale)
        self.view_features = view_features
        self.mask_features = mask_features
        # "mean" = lift_lseg_features (A4), "consensus" = lift_xdecoder_features (A5, the reference's default teacher),
        # "ape" = lift_ape_features (the APE teacher: the same consensus fusion on 256-d instance embeddings)
        self.lifting = lifting or ("consensus" if (mask_features is not None and view_features is None) else "mean")
        if self.lifting not in ("mean", "consensus", "ape"):
            raise ValueError(f"unknown lifting {self.lifting!r}")      # run/validation.py:106 raises ValueError too
        self.use_lseg = self.lifting == "mean"
        self.use_ape = self.lifting == "ape"                           # models/affinity_module.py:1496-1497
        if K is not None:
            self.K = K
        if n_pool is not None:
            self.n_pool = n_pool
        if sharpen is not None:
            self.affinity_sharpen_factor = sharpen
        self.min_visible = min_visible
        self.fill_cell = fill_cell
        self.last = {}
        self.stage_hook = None   # optional callable(name) invoked at stage boundaries (bench.py records CUDA events)

    def _mark(self, name):
        if self.stage_hook is not None:
            self.stage_hook(name)

    def eval(self):
        return self

    # ------------------------------------------------------------------ stage 1b (reference tuple form)
    @staticmethod
    def pixels_from_batch(batch_data, W):
        """x_labels / y_labels / mask_2ds of the 20-tuple -> pix [V,N] int32 on the device
        (dataset/data_loader_ablation.py:348-353, collate :446-459)."""
        scene_coords = batch_data[0]
        x_labels, y_labels, mask_2ds = batch_data[12], batch_data[13], batch_data[14]
        n = scene_coords.shape[0]
        v = mask_2ds.shape[0] // n
        mask = mask_2ds[:, 1].cuda().bool()
        pix = torch.full((v * n,), -1, dtype=torch.int32, device="cuda")
        lin = (x_labels.cuda().long() * W + y_labels.cuda().long()).to(torch.int32)
        if int(mask.sum()) != lin.shape[0]:
            raise L.GeoPurifyError("mask_2ds and x_labels disagree (a visible point with row or col 0 was dropped; "
                                   "the reference needs cut_boundary >= 1)")
        pix[mask] = lin
        return pix.view(v, n)

    @torch.no_grad()
    def lift_lseg_features(self, batch_data, max_batch_size=24):
        """Gather + mean over views + 1-NN fill (models/affinity_module.py:404-453).  `self.view_features(batch_data)`
        supplies the [V,H,W,D] channels-last maps the LSeg forward would have produced."""
        scene_coords = batch_data[0]
        feat = self.view_features(batch_data)
        v, h, w, d = feat.shape
        pix = self.pixels_from_batch(batch_data, w)
        out, seen = ops.lift_mean_pix(pix, None, feat, h, w)
        xyz = scene_coords.float().cuda().contiguous()
        nn = ops.nn1_fill(xyz, seen, self.fill_cell)
        ops.copy_rows(out, d, nn)
        return out, self.text_features, self.logit_scale

    @torch.no_grad()
    def lift_xdecoder_features(self, batch_data):
        """Mask-embedding gather, per-view fill, consensus top-3 fusion, scene-level 1-NN fill
        (models/affinity_module.py:455-714 minus the X-Decoder forward and its mask post-processing).
        `self.mask_features(batch_data)` supplies (idmap [V,H,W] uint8, table [V,Q,D] float32) on the device: the
        panoptic-style partition of :544-570 and the kept queries' embeddings."""
        idmap, table = self.mask_features(batch_data)
        return self._lift_masks(batch_data, idmap, table)

    @torch.no_grad()
    def lift_ape_features(self, batch_data):
        """APE teacher variant (models/affinity_module.py:736-993), feature_dim = 256 (:763).  Per view a point takes
        the MEAN of the embeddings of the final masks covering its pixel (:868-882); the final masks are the disjoint
        arg-max partition of :844-858, so at most one mask covers a pixel and the mean is that mask's embedding (or
        zero) -- the same gather as lift_xdecoder_features.  Per-view 1-NN fill (:895-911), normalise + logits
        (:913-916), consensus top-min(3, #views) softmax fusion (:945-969) and the scene-level fill (:971-981) are
        line for line those of the X-Decoder path, so both run through gp_lift_consensus.
        `self.mask_features(batch_data)` supplies (idmap [V,H,W] uint8, table [V,Q,256] float32): the partition and
        `outputs['scene_features_for_instances']` of the kept instances."""
        idmap, table = self.mask_features(batch_data)
        if table.shape[2] != self.feature_dim:
            raise L.GeoPurifyError(f"lift_ape_features: embeddings are {table.shape[2]}-d, model feature_dim is "
                                   f"{self.feature_dim} (the reference's APE path is 256-d)")
        return self._lift_masks(batch_data, idmap, table)

    def _lift_masks(self, batch_data, idmap, table):
        scene_coords = batch_data[0]
        w, d = idmap.shape[2], table.shape[2]
        pix = self.pixels_from_batch(batch_data, w)
        xyz = scene_coords.float().cuda().contiguous()
        out, seen = ops.lift_consensus(pix, None, idmap, table, self.text_features, self.logit_scale, xyz,
                                       self.fill_cell)
        nn = ops.nn1_fill(xyz, seen, self.fill_cell)
        ops.copy_rows(out, d, nn)
        return out, self.text_features, self.logit_scale

    # ------------------------------------------------------------------ stages 2-3
    @torch.no_grad()
    def purify(self, lifted, gauss, voxel_coords_i32, p2v_i32, extent, keep_intermediates=False):
        """Voxel scatter-mean, student, kNN, affinity softmax, pooling (models/affinity_module.py:1524-1589).
        Returns (pooled [M,D] in sorted voxel order, p2v_sorted [N] int32)."""
        with ops.scratch(not keep_intermediates):
            return self._purify(lifted, gauss, voxel_coords_i32, p2v_i32, extent, keep_intermediates)

    def _purify(self, lifted, gauss, voxel_coords_i32, p2v_i32, extent, keep_intermediates):
        d = self.feature_dim
        m = voxel_coords_i32.shape[0]
        index = ops.VoxelIndex(voxel_coords_i32, extent)
        p2v_sorted = index.rank[p2v_i32.long()].contiguous()
        plan = ops.ScatterPlan(p2v_sorted, m)
        stride = self.affinity_student.input_stride
        vin = ops._new((m, stride), torch.float32, "vin")
        if stride > d + 6:
            vin[:, d + 6:].zero_()       # only the pad columns: the two scatter-means write every other element
        plan.mean(lifted, d, vin, 0)
        plan.mean(gauss, 6, vin, d)
        self._mark("voxel_mean")
        nbr27 = index.kernel_map()
        embed = self.affinity_student.forward_sorted(vin, nbr27, normalize=True)
        self._mark("student")
        nbr_orig, nbr_sorted = index.knn(self.K, want_orig=keep_intermediates, want_sorted=True)
        self._mark("knn")
        w = ops.affinity_softmax(embed, nbr_sorted, self.affinity_sharpen_factor)
        self._mark("affinity")
        pooled = ops.pool(vin, d, nbr_sorted, w, self.n_pool)
        self._mark("pool")
        if keep_intermediates:
            self.last = dict(index=index, voxel_in=vin, embed=embed, knn_orig=nbr_orig, knn_sorted=nbr_sorted,
                             weights=w, pooled=pooled, p2v_sorted=p2v_sorted, nbr27=nbr27)
        return pooled, p2v_sorted

    @torch.no_grad()
    def evaluate_scene(self, batch_data, vis_prefix="scene0695_00", keep_intermediates=False):
        """Reference signature and return value (models/affinity_module.py:1490-1608)."""
        if self.use_lseg:
            f_full, text_features, logit_scale = self.lift_lseg_features(batch_data)      # :1495
        elif self.use_ape:
            f_full, text_features, logit_scale = self.lift_ape_features(batch_data)       # :1497
        else:
            f_full, text_features, logit_scale = self.lift_xdecoder_features(batch_data)  # :1500
        scene_coords_3d, inds = batch_data[1], batch_data[2]
        gauss = batch_data[19][:, :6].float().cuda().contiguous()
        vox = scene_coords_3d.to(torch.int32)
        vox = vox - vox.min(0).values          # any integer origin (MinkowskiEngine-style negative coordinates too):
        extent = tuple(int(e) + 1 for e in vox.max(0).values.tolist())   # kernel map and kNN are translation invariant
        if keep_intermediates:
            self.last_lifted = f_full
        pooled, p2v_sorted = self.purify(f_full, gauss, vox.cuda().contiguous(), inds.to(torch.int32).cuda(), extent,
                                         keep_intermediates)
        scene_features = ops.gather_rows(pooled, self.feature_dim, p2v_sorted)
        return {"scene_features": scene_features, "text_features": text_features, "logit_scale": logit_scale}

    # ------------------------------------------------------------------ B200-first whole-scene path
    @torch.no_grad()
    def run_scene(self, s: SceneInputs, hist=None, ignore=None, keep_intermediates=False, want_pred=False):
        """Stages 1a-4 on the device: fused project + gather-mean, fill, purify, per-voxel classification,
        label gather + IoU histograms.  Returns dict(labels_voxel, p2v_sorted, hist, seen)."""
        with ops.scratch(not keep_intermediates):
            return self._run_scene(s, hist, ignore, keep_intermediates, want_pred)

    def _run_scene(self, s, hist, ignore, keep_intermediates, want_pred):
        W, H = s.image_dim
        d = self.feature_dim
        self._mark("begin")
        pix, counts = ops.project_views(s.coords, s.w2c, s.intr, s.depth, H, W, s.cut, s.vis_thres)
        keep = ops.accept_views(counts, self.min_visible)
        self._mark("project")
        if self.lifting in ("consensus", "ape"):
            lifted, seen = ops.lift_consensus(pix, keep, s.idmaps, s.tables, self.text_features, self.logit_scale,
                                              s.coords_f32, self.fill_cell)
        elif s.feat is None:      # compact 2-D form only: gather through the id map, no dense map
            lifted, seen = ops.lift_mean_idmap(s.coords, s.w2c, s.intr, s.depth, H, W, s.cut, s.vis_thres, keep,
                                               s.idmaps, s.tables)
        else:
            lifted, seen = ops.lift_mean_fused(s.coords, s.w2c, s.intr, s.depth, H, W, s.cut, s.vis_thres, keep,
                                               s.feat)
        self._mark("lift")
        nn = ops.nn1_fill(s.coords_f32, seen, self.fill_cell)
        ops.copy_rows(lifted, d, nn)
        self._mark("fill")
        if keep_intermediates:
            self.last_lifted, self.last_seen, self.last_keep, self.last_counts = lifted, seen, keep, counts
        pooled, p2v_sorted = self.purify(lifted, s.gauss, s.voxel_coords, s.p2v, s.extent, keep_intermediates)
        labels_vox, _, _ = ops.logits_argmax(pooled, d, self.text_features, self.logit_scale, want_zero=False)
        out = dict(labels_voxel=labels_vox, p2v_sorted=p2v_sorted, seen=seen, pooled=pooled)
        if s.labels is not None:
            c = self.text_features.shape[0]
            out["hist"], out["pred"] = ops.iou_hist(labels_vox, p2v_sorted, s.labels, ignore, c, hist,
                                                    want_pred=keep_intermediates or want_pred)
        self._mark("classify")
        return out


GeoPurifyB200 = SonataXAffinityTrainer


class InterleavedScenes:
    """Runs a stream of independent device-resident scenes on `n_streams` CUDA streams, scene i on stream i mod n, so
    that the short latency-bound stages of one scene (index build, sorts, fill, projection) can run beside the
    persistent tensor-core kernels of another.  Results are identical to calling `run_scene` per scene: each scene's
    tensors live on its own stream (the caching allocator is stream-aware) and the shared histogram block is only
    touched by int64 atomics."""

    def __init__(self, model, n_streams=2):
        self.model = model
        self.streams = [torch.cuda.Stream() for _ in range(max(1, int(n_streams)))]

    def run(self, scenes, hist=None, ignore=None, collect=True):
        """`collect=False` drops every scene's result dict at once (an evaluation that only wants the accumulated
        `hist` would otherwise keep one [M,512] pooled tensor per scene alive)."""
        main = torch.cuda.current_stream()
        for st in self.streams:
            st.wait_stream(main)
        outs = []
        for i, s in enumerate(scenes):
            with torch.cuda.stream(self.streams[i % len(self.streams)]):
                r = self.model.run_scene(s, hist=hist, ignore=ignore)
                if collect:
                    outs.append(r)
                del r
        for st in self.streams:
            main.wait_stream(st)
        return outs


def scene_inputs_from_synth(scene, feat_dtype=torch.bfloat16, device="cuda", coords_f64=True):
    """Upload a `geopurify_b200.synth.Scene` (host numpy) to the device in the layout `run_scene` expects.
    The dense feature map is the table lookup of the id map -- the stand-in for the 2-D VLM output."""
    from .synth import mapper_intrinsics
    dev = torch.device(device)
    W, H = scene.image_dim
    k = mapper_intrinsics(scene.intrinsics, scene.image_dim)
    intr = np.array([k[0, 0], k[1, 1], k[0, 2], k[1, 2]], dtype=np.float64)
    w2c, intr_d = ops.cameras_to_device(scene.w2c, intr)
    c32 = torch.as_tensor(scene.coords).to(dev)
    coords = c32.double() if coords_f64 else c32
    tables = torch.as_tensor(scene.tables).to(dev).to(feat_dtype)
    idm = torch.as_tensor(scene.idmaps).to(dev).long()
    feat = torch.stack([tables[v][idm[v]] for v in range(scene.n_views)]).contiguous()
    vox = torch.as_tensor(scene.voxel_coords.astype(np.int32)).to(dev)
    extent = tuple(int(e) + 1 for e in scene.voxel_coords.max(0))
    return SceneInputs(coords=coords.contiguous(), coords_f32=c32.contiguous(),
                       gauss=torch.as_tensor(scene.gauss).to(dev).contiguous(), voxel_coords=vox.contiguous(),
                       p2v=torch.as_tensor(scene.inds_reconstruct.astype(np.int32)).to(dev), extent=extent, w2c=w2c,
                       intr=intr_d, depth=torch.as_tensor(scene.depth).to(dev).contiguous(), feat=feat,
                       image_dim=(W, H), cut=scene.cut, vis_thres=scene.vis_thres,
                       labels=torch.as_tensor(scene.labels).to(dev),
                       idmaps=torch.as_tensor(scene.idmaps).to(dev).contiguous(),
                       tables=torch.as_tensor(scene.tables).to(dev).float().contiguous())


# ---------------------------------------------------------------------- host-buffer entry (the e2e call)
HOST_FIELDS = ("coords", "coords_f32", "gauss", "voxel_coords", "p2v", "w2c", "intr", "depth", "idmaps", "tables",
               "labels")


@dataclass
class HostScene:
    """One scene as pinned HOST tensors -- what a caller holds after the data loader and the 2-D VLM
    (the VLM output in its compact (id map, embedding table) form).  `nbytes` is what one upload moves."""
    t: dict
    extent: tuple
    image_dim: tuple
    cut: int
    vis_thres: float

    @property
    def nbytes(self):
        return sum(v.numel() * v.element_size() for v in self.t.values())

    @property
    def n_points(self):
        return self.t["coords"].shape[0]


def host_scene_from_synth(scene, pin=True):
    from .synth import mapper_intrinsics
    k = mapper_intrinsics(scene.intrinsics, scene.image_dim)
    V = scene.n_views
    intr = np.repeat(np.array([[k[0, 0], k[1, 1], k[0, 2], k[1, 2]]], dtype=np.float64), V, 0)
    t = dict(coords=torch.as_tensor(scene.coords.astype(np.float64)),
             coords_f32=torch.as_tensor(scene.coords),
             gauss=torch.as_tensor(scene.gauss),
             voxel_coords=torch.as_tensor(scene.voxel_coords.astype(np.int32)),
             p2v=torch.as_tensor(scene.inds_reconstruct.astype(np.int32)),
             w2c=torch.as_tensor(scene.w2c.astype(np.float64).reshape(V, 16)),
             intr=torch.as_tensor(intr),
             depth=torch.as_tensor(scene.depth),
             idmaps=torch.as_tensor(scene.idmaps),
             tables=torch.as_tensor(scene.tables),
             labels=torch.as_tensor(scene.labels))
    t = {n: (v.contiguous().pin_memory() if pin else v.contiguous()) for n, v in t.items()}
    extent = tuple(int(e) + 1 for e in scene.voxel_coords.max(0))
    return HostScene(t=t, extent=extent, image_dim=tuple(scene.image_dim), cut=scene.cut, vis_thres=scene.vis_thres)


def upload_scene(h: HostScene, feat_dtype=torch.bfloat16, feat_out=None, dense=True):
    """Host -> device copies of every field (async on the current stream, pinned sources) + the dense feature map
    built on the device from (idmaps, tables) (`dense=False` skips it: consensus lifting reads the compact form).
    Returns SceneInputs."""
    d = {n: v.cuda(non_blocking=True) for n, v in h.t.items()}
    feat = ops.expand_idmap(d["idmaps"], d["tables"], feat_dtype, out=feat_out) if dense else None
    return SceneInputs(coords=d["coords"], coords_f32=d["coords_f32"], gauss=d["gauss"],
                       voxel_coords=d["voxel_coords"], p2v=d["p2v"], extent=h.extent, w2c=d["w2c"], intr=d["intr"],
                       depth=d["depth"], feat=feat, image_dim=h.image_dim, cut=h.cut, vis_thres=h.vis_thres,
                       labels=d["labels"], idmaps=d["idmaps"], tables=d["tables"])


def run_scene_host(model, h: HostScene, ignore, feat_dtype=torch.bfloat16, feat_out=None):
    """End-to-end call on HOST buffers: upload, purify + classify, read back per-point predictions and the
    IoU histogram block.  Returns (pred [N] int64 CPU, hist [3,C] int64 CPU)."""
    s = upload_scene(h, feat_dtype, feat_out)
    out = model.run_scene(s, ignore=ignore, want_pred=True)
    return out["pred"].cpu(), out["hist"].cpu()


class HostScenePipeline:
    """`run_scene_host` for a stream of scenes with the copies taken off the critical path: scene i+1 is uploaded on a
    side stream while scene i runs; results come back through pinned buffers with a second side stream.  Every scene
    still pays its full host->device and device->host traffic -- the copies overlap the previous / next scene's
    kernels instead of serialising with them.  Results are identical to calling `run_scene_host` per scene."""

    def __init__(self, model, ignore, feat_dtype=torch.bfloat16, dense=None, n_streams=2):
        self.model, self.ignore, self.feat_dtype = model, ignore, feat_dtype
        # scenes alternate between `n_streams` compute streams (as InterleavedScenes): the short stages of one scene run
        # beside the tensor-core kernels of the other
        self.compute = [torch.cuda.Stream() for _ in range(max(1, int(n_streams)))]
        # the compact (id map, table) form is lifted directly; dense=True expands it to [V,H,W,D] maps first
        self.dense = False if dense is None else dense
        self.copy_in = torch.cuda.Stream()
        self.copy_out = torch.cuda.Stream()
        self._feat = [None, None]     # two dense feature buffers, reused (they are the largest allocation)
        self._pinned = {}             # result buffers in pinned host memory, grown on demand (cudaHostAlloc is slow)

    def _host_out(self, key, shape, dtype):
        n = int(np.prod(shape))
        buf = self._pinned.get(key)
        if buf is None or buf.numel() < n or buf.dtype != dtype:
            buf = torch.empty(max(n, 1), dtype=dtype, pin_memory=True)
            self._pinned[key] = buf
        return buf[:n].view(shape)

    def _upload(self, h, slot):
        with torch.cuda.stream(self.copy_in):
            for st in self.compute:                   # buffers of slot are free once earlier compute work is done
                self.copy_in.wait_stream(st)
            fo = self._feat[slot]
            if self.dense and fo is not None:
                V, (W, H) = h.t["idmaps"].shape[0], h.image_dim
                if fo.shape[0] < V or fo.shape[1] != H or fo.shape[2] != W:
                    fo = None
                else:
                    fo = fo[:V]
            s = upload_scene(h, self.feat_dtype, feat_out=fo, dense=self.dense)
            if self.dense and (self._feat[slot] is None or self._feat[slot].shape[0] < s.feat.shape[0]):
                self._feat[slot] = s.feat
            ev = torch.cuda.Event()
            ev.record(self.copy_in)
        return s, ev

    def run(self, hosts):
        """hosts: sequence of HostScene.  Returns [(pred [N] int64 pinned CPU, hist [3,C] int64 pinned CPU), ...]."""
        hosts = list(hosts)
        out = []
        caller = torch.cuda.current_stream()
        for st in self.compute:
            st.wait_stream(caller)
        nxt = self._upload(hosts[0], 0) if hosts else None
        for i, h in enumerate(hosts):
            s, ev = nxt
            comp = self.compute[i % len(self.compute)]
            comp.wait_event(ev)
            for t in (s.coords, s.feat, s.depth, s.idmaps, s.tables, s.labels, s.voxel_coords, s.p2v, s.gauss,
                      s.coords_f32, s.w2c, s.intr):
                if t is not None:
                    t.record_stream(comp)
            if i + 1 < len(hosts):
                nxt = self._upload(hosts[i + 1], (i + 1) & 1)
            with torch.cuda.stream(comp):
                res = self.model.run_scene(s, ignore=self.ignore, want_pred=True)
                done = torch.cuda.Event()
                done.record(comp)
            with torch.cuda.stream(self.copy_out):
                self.copy_out.wait_event(done)
                # one pinned staging buffer per scene of the batch, reused by later run() calls
                pred = self._host_out(("pred", i), res["pred"].shape, res["pred"].dtype)
                hist = self._host_out(("hist", i), res["hist"].shape, res["hist"].dtype)
                pred.copy_(res["pred"], non_blocking=True)
                hist.copy_(res["hist"], non_blocking=True)
                res["pred"].record_stream(self.copy_out)
                res["hist"].record_stream(self.copy_out)
            out.append((pred, hist))
        self.copy_out.synchronize()
        for st in self.compute:
            caller.wait_stream(st)
        return out
