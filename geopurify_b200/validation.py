"""Mirror of the reference's classification + metric code on the CUDA path.
(`AverageMeter` is the reference's interface class, util/util.py:108-124, kept verbatim so callers can swap imports.)

Reference: run/validation.py:413-439 (normalise, logits, argmax, zero-row label fill, IoU call) and
:458-553 (running sums, mIoU / mAcc / allAcc); util/util.py:108-124 (AverageMeter), :160-177
(intersectionAndUnionGPU).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L
from . import ops


class AverageMeter:
    def __init__(self):
        self.reset()

    def reset(self):
        self.val = 0
        self.avg = 0
        self.sum = 0
        self.count = 0

    def update(self, val, n=1):
        self.val = val
        self.sum += val * n
        self.count += n
        self.avg = self.sum / self.count


def intersectionAndUnionGPU(output, target, K, ignore_indexs=[255]):
    """util/util.py:160-177: `output[target == ign] = ign` IN PLACE, then three K-bin histograms as float32
    CUDA tensors (area_intersection, area_union, area_target)."""
    L.require_cuda()
    assert output.dim() in [1, 2, 3, 4]
    assert output.shape == target.shape
    out_flat = output.view(-1)
    tgt = target.view(-1).to(torch.int64).cuda().contiguous()
    lab = out_flat.to(torch.int32).cuda().contiguous()
    ign = torch.as_tensor(list(ignore_indexs), dtype=torch.int64).cuda()
    hist, pred = ops.iou_hist(lab, None, tgt, ign, K, want_pred=True)
    out_flat.copy_(pred.to(out_flat.dtype).to(out_flat.device))
    h = hist.to(torch.float32)
    return h[0], h[1], h[2]


@torch.no_grad()
def classify_scene(scene_features, text_features, logit_scale, scene_coords=None, fill_cell=0.05):
    """run/validation.py:413-432: returns `scene_pred_2d` [N] int64 on the device.  Rows whose normalised
    features are all zero copy the label of the nearest non-zero row, measured on columns 1:4 of the [N,3]
    coordinate array, i.e. on (y, z) only -- the reference's quirk, kept."""
    d = scene_features.shape[1]
    label, zero, _ = ops.logits_argmax(scene_features.contiguous(), d, text_features.float().cuda().contiguous(),
                                       float(logit_scale), want_zero=True)
    if scene_coords is not None and bool(zero.any()) and not bool(zero.all()):
        yz = scene_coords.float().cuda().clone()
        yz[:, 0] = 0.0
        nn = ops.nn1_fill(yz.contiguous(), (1 - zero).to(torch.uint8), fill_cell)
        label = label[nn.long()]
    return label.long()


def mean_iou(inter_sum, union_sum, target_sum):
    """run/validation.py:513-523 on float32 numpy sums."""
    i = np.asarray(inter_sum, dtype=np.float32)
    u = np.asarray(union_sum, dtype=np.float32)
    t = np.asarray(target_sum, dtype=np.float32)
    iou_class = i / (u + 1e-10)
    acc_class = i / (t + 1e-10)
    return float(np.mean(iou_class)), float(np.mean(acc_class)), float(sum(i) / (sum(t) + 1e-10))


def shard_scenes(n_scenes, rank=0, world=1, sizes=None):
    """Scene indices of `rank` out of `world` processes (one process per GPU).
    Default: round-robin i = r (mod W).  With `sizes` (points per scene) scenes are dealt longest-first so the
    ranks finish together.  The reference's own contiguous split is `get_batch_scenes` (run/validation.py:269-280)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    order = list(range(n_scenes))
    if sizes is not None:
        if len(sizes) != n_scenes:
            raise ValueError("sizes must have one entry per scene")
        order.sort(key=lambda i: (-int(sizes[i]), i))
    return order[rank::world]


def reduce_histograms(hist, group=None):
    """The path's one exchange step: sum the [3,C] int64 histogram block over all ranks (NCCL on CUDA tensors,
    gloo on CPU tensors).  Replaces the dead all_reduce calls at run/validation.py:441-450.  No-op when
    torch.distributed is not initialised."""
    import torch.distributed as dist
    if hist.dtype != torch.int64:
        raise TypeError("histograms are accumulated in int64 (exact) and reduced as such")
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)
    return hist


class ValidateResult(tuple):
    """What the reference's validate() returns -- `(mIoU_Base, mIoU_Novel)` (run/validation.py:549-552) -- with the
    numbers it only logs attached: `.base`, `.novel`, `.all` = (mIoU, mAcc, allAcc) triples, `.iou_class_*` and the
    exact int64 running sums `.sums` = (intersection, union, target) over all classes."""

    def __new__(cls, base, novel, all_, iou_class, sums):
        self = super().__new__(cls, (base[0], novel[0]))
        self.base, self.novel, self.all = base, novel, all_
        self.iou_class_base, self.iou_class_novel, self.iou_class_all = iou_class
        self.sums = sums
        return self


def _split_indices(category_split, name):
    idx = category_split[name] if isinstance(category_split, dict) else getattr(category_split, name)
    return np.asarray(list(idx), dtype=np.int64)


@torch.no_grad()
def validate(val_loader, model, test_classes, test_ignore_label, logger=None, rank=0, world=1, category_split=None):
    """Loop body of run/validation.py:376-553.  The three [C] histograms of every scene are accumulated in int64
    (exact; the reference's float32 meters are exact only below 2^24 points per class, and identical there) and the
    Base / Novel / All metrics of :478-519 are taken from the float32 view of the sums, as the reference's numpy
    meters would hold them.  With world > 1 each rank evaluates the scenes `shard_scenes` gives it and the int64
    block is all-reduced once at the end, so every rank returns the same numbers.

    Without `category_split`: returns ((mIoU, mAcc, allAcc) over all classes, (intersection, union, target) float32).
    With `category_split` (an object or dict with `base_category` / `novel_category` index lists, the reference's
    `args.category_split`): returns a `ValidateResult`, i.e. the reference's `(mIoU_Base, mIoU_Novel)` tuple."""
    c = test_classes
    acc = np.zeros((3, c), dtype=np.int64)
    mine = set(shard_scenes(len(val_loader), rank, world)) if world > 1 else None
    for i, batch_data in enumerate(val_loader):
        if mine is not None and i not in mine:
            continue
        if batch_data is None:                                         # run/validation.py:377-379
            continue
        scene_coords, scene_label = batch_data[0], batch_data[3]
        res = model.evaluate_scene(batch_data)
        pred = classify_scene(res["scene_features"], res["text_features"], res["logit_scale"], scene_coords)
        inter, union, tgt = intersectionAndUnionGPU(pred, scene_label.detach().cuda(), test_classes, test_ignore_label)
        acc += torch.stack([inter, union, tgt]).round().to(torch.int64).cpu().numpy()   # per-scene counts < 2^24: exact
        if logger:
            logger.info("Process: [{}/{}]".format(i, len(val_loader)))
    if world > 1:
        block = torch.as_tensor(acc)
        if torch.distributed.get_backend() == "nccl":
            block = block.cuda()
        acc = reduce_histograms(block).cpu().numpy()
    sums = [acc[j].astype(np.float32) for j in range(3)]
    all_ = mean_iou(sums[0], sums[1], sums[2])
    if category_split is None:
        return all_, (sums[0], sums[1], sums[2])
    base_i, novel_i = _split_indices(category_split, "base_category"), _split_indices(category_split, "novel_category")
    base = mean_iou(sums[0][base_i], sums[1][base_i], sums[2][base_i])
    novel = mean_iou(sums[0][novel_i], sums[1][novel_i], sums[2][novel_i])
    iou_class = sums[0] / (sums[1] + np.float32(1e-10))
    if logger:
        logger.info("Val 2d result: mIoU_Base/mAcc_Base/allAcc_Base {:.4f}/{:.4f}/{:.4f}.".format(*base))
        logger.info("Val 2d result: mIoU_Novel/mAcc_Novel/allAcc_Novel {:.4f}/{:.4f}/{:.4f}.".format(*novel))
        logger.info("Val 2d result: mIoU_All/mAcc_All/allAcc_All {:.4f}/{:.4f}/{:.4f}.".format(*all_))
    return ValidateResult(base, novel, all_, (iou_class[base_i], iou_class[novel_i], iou_class), tuple(acc))
