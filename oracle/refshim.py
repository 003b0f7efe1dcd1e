"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.

Imports the *real* reference (read-only at /root/reference) in an environment
without detectron2 / pytorch3d / matplotlib / segment_anything, so that its own
functions can be executed to (a) generate tests/golden/ fixtures
(tests/golden/make_golden.py) and (b) validate the oracle restatements live when
/root/reference is present (authoring container only; it does not exist on the
GPU box and nothing in -m gpu tests, smoke() or bench.py touches it).

Recipe: a sys.meta_path finder fabricates permissive stub
packages for the missing third-party prefixes; real stand-ins from
oracle/standins.py replace the handful of third-party functions whose
arithmetic the hot path depends on; roi_heads.py is loaded by file path because
the snapshot's package __init__ imports a ``build_cube_head`` that does not
exist (roi_heads.py:31 vs cube_head.py:1-197).
"""
from __future__ import annotations

import importlib
import importlib.abc
import importlib.machinery
import importlib.util
import os
import sys
import types

REF_ROOT = os.environ.get("WCR_REFERENCE_ROOT", "/root/reference")

_STUB_PREFIXES = (
    "detectron2", "pytorch3d", "matplotlib", "segment_anything", "fvcore",
    "iopath", "pycocotools", "wandb", "groundingdino", "GroundingDINO",
    "pyransac3d", "open3d", "cv2", "seaborn", "termcolor", "tabulate",
)


class _Dummy:
    """Permissive placeholder: callable, subclassable, attribute-chaining."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        # used as decorator (``@configurable`` / ``@REGISTRY.register()``)
        if len(a) == 1 and callable(a[0]) and not k:
            return a[0]
        return _Dummy()

    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        return _Dummy()

    def __mro_entries__(self, bases):
        return (object,)

    def __iter__(self):
        return iter(())

    def register(self, *a, **k):
        if len(a) == 1 and callable(a[0]):
            return a[0]
        return lambda f: f


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        return _Dummy()


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        top = fullname.split(".")[0]
        if top in _STUB_PREFIXES:
            # prefer a real install if there is one (e.g. cv2 is present here)
            if top in ("cv2", "tabulate", "termcolor"):
                return None
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        m = _StubModule(spec.name)
        m.__path__ = []
        return m

    def exec_module(self, module):
        pass


_installed = False


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "ProposalNetwork", "utils"))


def _install():
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError(f"reference not present at {REF_ROOT}")
    from . import standins

    sys.meta_path.append(_StubFinder())
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)

    # real stand-ins where arithmetic matters
    d2s = importlib.import_module("detectron2.structures")
    d2s.Boxes = standins.Boxes
    d2s.pairwise_iou = standins.pairwise_iou
    ev = importlib.import_module("detectron2.utils.events")
    ev.get_event_storage = standins.get_event_storage
    p3t = importlib.import_module("pytorch3d.transforms")
    p3t.axis_angle_to_matrix = standins.axis_angle_to_matrix
    p3t.rotation_6d_to_matrix = standins.rotation_6d_to_matrix
    p3so3 = importlib.import_module("pytorch3d.transforms.so3")
    p3so3.so3_relative_angle = standins.so3_relative_angle
    tvo = importlib.import_module("torchvision.ops")   # real torchvision: keep its focal / GIoU losses
    assert hasattr(tvo, "sigmoid_focal_loss")
    _installed = True


def _load_by_path(modname: str, relpath: str):
    spec = importlib.util.spec_from_file_location(modname, os.path.join(REF_ROOT, relpath))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[modname] = mod
    spec.loader.exec_module(mod)
    return mod


_cache: dict = {}


def load():
    """Return a namespace with the reference's hot-path modules.

    .plane .spaces .conversions .utils  -> ProposalNetwork/utils/*.py
    .math_util                          -> cubercnn/util/math_util.py
    .roi_heads                          -> cubercnn/modeling/roi_heads/roi_heads.py
    """
    if "ns" in _cache:
        return _cache["ns"]
    _install()
    ns = types.SimpleNamespace()
    ns.plane = _load_by_path("_wcr_ref_plane", "ProposalNetwork/utils/plane.py")
    ns.math_util = importlib.import_module("cubercnn.util.math_util")
    ns.spaces = importlib.import_module("ProposalNetwork.utils.spaces")
    ns.conversions = importlib.import_module("ProposalNetwork.utils.conversions")
    ns.utils = importlib.import_module("ProposalNetwork.utils.utils")

    # roi_heads: hand-made package, inject the missing build_cube_head first
    pkg_name = "cubercnn.modeling.roi_heads"
    if pkg_name not in sys.modules or not hasattr(sys.modules[pkg_name], "_wcr_handmade"):
        for parent in ("cubercnn.modeling",):
            if parent not in sys.modules:
                pm = types.ModuleType(parent)
                pm.__path__ = [os.path.join(REF_ROOT, *parent.split("."))]
                sys.modules[parent] = pm
        pkg = types.ModuleType(pkg_name)
        pkg.__path__ = [os.path.join(REF_ROOT, "cubercnn/modeling/roi_heads")]
        pkg._wcr_handmade = True
        sys.modules[pkg_name] = pkg
        ch = _load_by_path(pkg_name + ".cube_head", "cubercnn/modeling/roi_heads/cube_head.py")
        if not hasattr(ch, "build_cube_head"):
            ch.build_cube_head = lambda cfg, shape: ch.ROI_CUBE_HEAD_REGISTRY.get(
                cfg.MODEL.ROI_CUBE_HEAD.NAME)(cfg, shape)
        # first-party modules roi_heads.py imports but the hot path never calls
        for stub in ("cubercnn.data.generate_ground_segmentations",
                     "cubercnn.modeling.proposal_generator.rpn",
                     "cubercnn.modeling.roi_heads.fast_rcnn",
                     "cubercnn.data", "cubercnn.modeling.proposal_generator"):
            if stub not in sys.modules:
                sm = _StubModule(stub)
                sm.__path__ = []
                sys.modules[stub] = sm
        ns.roi_heads = _load_by_path(pkg_name + ".roi_heads",
                                     "cubercnn/modeling/roi_heads/roi_heads.py")
    else:
        ns.roi_heads = sys.modules[pkg_name + ".roi_heads"]
    _cache["ns"] = ns
    return ns


def roi_head_instance():
    """A bare ROIHeads3DScore whose loss helper methods can be called unbound."""
    ns = load()
    obj = object.__new__(ns.roi_heads.ROIHeads3DScore)
    return obj
