import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def small_scene():
    from geopurify_b200 import synth
    return synth.make_scene(n_points=12000, n_views=6, image_dim=(80, 60), seed=3, room=(2.0, 1.6, 2.0), n_boxes=2,
                            cut=2, q_range=(4, 12), feat_dim=128, n_classes=7)


@pytest.fixture(scope="session")
def small_state():
    from geopurify_b200 import synth
    return synth.make_student_state(input_dim=134, hidden_dim=128, embed_dim=128, seed=3)


@pytest.fixture(autouse=True)
def _workspace_guards(request):
    """GEOPURIFY_B200_GUARD=1 python -m pytest tests -m gpu: every workspace of every GPU test gets 4 KB canaries on both
    sides (ops._ws), verified after each test -- the memcheck substitute on a pool where compute-sanitizer is closed."""
    yield
    if os.environ.get("GEOPURIFY_B200_GUARD", "0") == "1" and request.node.get_closest_marker("gpu"):
        from geopurify_b200 import ops
        ops.check_guards()
