"""CPU suite: the C-ABI library loads and exports every symbol include/geopurify_b200.h declares; host-side
mirrors keep the reference's names; product code does not import the oracle; no compute without a GPU."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

import geopurify_b200
from geopurify_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "geopurify_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gp_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert sorted(_lib.EXPORTS) == names, "python binding and header disagree"
    assert _lib.load().gp_version() >= 100


def test_argument_errors_need_no_gpu():
    lib = _lib.load()
    rc = lib.gp_knn_grid(None, 10, 8, 8, 8, 96, None, None, None)       # m < K+1
    assert rc == 1 and b"K+1" in lib.gp_last_error()
    rc = lib.gp_knn_grid(None, 1000, 8, 8, 8, 500, None, None, None)    # K > 128
    assert rc == 1
    assert lib.gp_voxel_index_workspace_bytes(1000, 100000, 8, 8) == 0  # extent too large
    assert lib.gp_voxel_index_workspace_bytes(1000, 600, 500, 150) > 0


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "geopurify_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            text = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in text.replace("the oracle", ""), fn


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_fails_loudly_without_gpu():
    from geopurify_b200.fusion_util import PointCloudToImageMapper
    from geopurify_b200 import synth
    m = PointCloudToImageMapper([320, 240], 0.05, 10, synth.scannet_intrinsics())
    with pytest.raises(geopurify_b200.GeoPurifyError):
        m.compute_mapping(np.eye(4, dtype=np.float32), np.zeros((4, 3)), None)


def test_student_state_dict_uses_reference_names():
    from geopurify_b200.student import AffinityPredictor, load_numpy_state
    from geopurify_b200 import synth
    model = AffinityPredictor(input_dim=518, embed_dim=128, hidden_dim=512)
    keys = set(model.state_dict())
    for k in ["input_layer.0.kernel", "input_layer.1.bn.weight", "input_layer.1.bn.running_var",
              "res_blocks.0.conv1.kernel", "res_blocks.3.norm2.bn.bias", "output_layer.kernel"]:
        assert k in keys
    assert model.state_dict()["input_layer.0.kernel"].shape == (27, 518, 512)
    assert model.state_dict()["output_layer.kernel"].shape == (512, 128)
    assert set(model.get_param_groups()) == {"input", "middle", "output"}
    small = AffinityPredictor(134, 128, 128)
    load_numpy_state(small, synth.make_student_state(134, 128, 128, seed=1))
    desc, flat = small.pack(torch.device("cpu"))
    assert desc.n_layers == 10 and desc.cin[0] == 160 and desc.residual_from[2] == 0 and desc.residual_from[1] == -1
    assert desc.kv[9] == 1 and desc.bn[9] == 0 and flat.numel() > 27 * 160 * 128


def test_mapper_intrinsics_rescale_matches_reference_ctor():
    from geopurify_b200.fusion_util import PointCloudToImageMapper
    from geopurify_b200 import synth
    from oracle import oracle
    intr = synth.scannet_intrinsics()
    m = PointCloudToImageMapper([320, 240], 0.05, 10, intr)
    assert np.array_equal(m.intrinsics, oracle.mapper_intrinsics(intr, (320, 240)))
    assert m.intrinsics[0, 2] == 160 and m.intrinsics[1, 2] == 120


def test_intrinsic_helpers_match_reference_golden():
    """make_intrinsic / adjust_intrinsic (models/utils/fusion_util.py:7-33) against the reference's own outputs."""
    import json
    import numpy as np
    from geopurify_b200 import fusion_util, synth
    cases = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "intrinsics.json")))
    for c in cases:
        for mod in (fusion_util, synth):
            k = mod.make_intrinsic(*c["args"])
            assert (np.asarray(k) == np.asarray(c["make"])).all()
            a = mod.adjust_intrinsic(np.asarray(k).copy(), c["src"], c["dst"])
            assert (np.asarray(a) == np.asarray(c["adjust"])).all()


def test_static_scratch_entry_point_limits_without_a_gpu():
    """gp_scatter_mean_static rejects sizes beyond its stated maximum before it touches CUDA, and the limits the GPU
    test file assumes are the header's."""
    src = open(os.path.join(ROOT, "include", "geopurify_b200.h")).read()
    lim = {k: int(v) for k, v in re.findall(r"#define (GP_SCATTER_STATIC_MAX_[NM]) (\d+)", src)}
    assert lim == {"GP_SCATTER_STATIC_MAX_N": 262144, "GP_SCATTER_STATIC_MAX_M": 262144}
    assert re.search(r"#define GP_SCATTER_STATIC_BYTES \(16 \* 1024 \* 1024\)", src)
    zz = open(os.path.join(ROOT, "tests", "test_zz_no_alloc.py")).read()
    assert "MAX_N = 262144" in zz and "MAX_M = 262144" in zz
    lib = _lib.load()
    for n, m in ((lim["GP_SCATTER_STATIC_MAX_N"] + 1, 8), (8, lim["GP_SCATTER_STATIC_MAX_M"] + 1)):
        rc = lib.gp_scatter_mean_static(None, n, m, None, 4, 4, None, 4, 0, None)
        assert rc == 1 and b"beyond the static scratch limits" in lib.gp_last_error()
    rc = lib.gp_scatter_mean_static(None, 8, 8, None, 2, 4, None, 4, 0, None)   # src_stride < D
    assert rc == 1 and b"bad sizes" in lib.gp_last_error()
