"""Stand-ins that let the reference's own inference code (`models/affinity_module.py:evaluate_scene`,
`lift_lseg_features`, `dataset/data_loader_ablation.py:scene_based_collate_fn`) execute in a container that lacks its
third-party dependencies.  TEST INFRASTRUCTURE: used only by `tests/golden/make_golden_scene.py` to generate golden
vectors; nothing under `geopurify_b200/` imports it.

Two kinds of stand-in:
* operators ON the path get a definition stub written from the library's published semantics
  (`me_stub` = MinkowskiEngine 0.5 subset, `faiss_stub` = IndexFlatL2, `scatter_stub` = torch_scatter.scatter_mean);
* libraries the path never calls (clip, sonata, open3d, xdecoder, detectron2, matplotlib, cv2, ...) become empty
  auto-attribute modules so that the reference's module-level imports succeed.
"""
import importlib
import importlib.abc
import importlib.machinery
import sys
import types

from . import faiss_stub, me_stub, scatter_stub

EMPTY = ("clip", "sonata", "open3d", "xdecoder", "detectron2", "matplotlib", "mpl_toolkits", "seaborn", "plotly",
         "cv2", "imageio", "trimesh", "plyfile", "tensorboardX", "SharedArray", "wandb", "timm", "ftfy", "encoding",
         "pytorch_lightning", "torchvision", "PIL", "skimage", "pointops", "spconv", "addict", "huggingface_hub",
         "omegaconf", "hydra", "kornia", "nltk", "pycocotools", "shapely", "fvcore", "iopath", "diff_gaussian_rasterization",
         "simple_knn", "plyfile", "lpips", "pytorch3d", "termcolor", "easydict")
stubbed = []


class _Auto(types.ModuleType):
    """Module whose every attribute is an empty class (enough for `from x import Y` and for base classes)."""

    def __getattr__(self, key):
        if key.startswith("__"):
            raise AttributeError(key)
        sub = self.__name__ + "." + key
        if sub in sys.modules:
            return sys.modules[sub]
        return type(key, (), {})


class _Finder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, name, path=None, target=None):
        if name.split(".")[0] in EMPTY:
            return importlib.machinery.ModuleSpec(name, self, is_package=True)
        return None

    def create_module(self, spec):
        m = _Auto(spec.name)
        m.__path__ = []
        stubbed.append(spec.name)
        return m

    def exec_module(self, module):
        pass


def install():
    """Register the stand-ins.  Real installed packages win: the finder sits at the END of sys.meta_path."""
    sys.modules.setdefault("MinkowskiEngine", me_stub)
    sys.modules.setdefault("MinkowskiEngine.MinkowskiFunctional", me_stub.MinkowskiFunctional)
    sys.modules.setdefault("torch_scatter", scatter_stub)
    sys.modules.setdefault("faiss", faiss_stub)
    if not any(isinstance(f, _Finder) for f in sys.meta_path):
        sys.meta_path.append(_Finder())
