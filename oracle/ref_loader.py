"""Import the two reference source files verbatim, without mmcv / mmdet.

Only usable where ``/root/reference`` exists (the build container).  It is
used by ``tests/golden/make_golden.py`` to produce the committed fixtures and
by the ``-m "not gpu"`` tests that re-validate the oracle against the live
reference when it is present; nothing that runs on the GPU box touches it.

The reference needs exactly five third-party names:
``mmdet.models.{DETECTORS,NECKS,build_backbone,build_head,build_neck}``,
``mmdet.models.detectors.BaseDetector``, ``mmcv.runner.auto_fp16`` and
``mmdet3d.core.bbox3d2result``.  They are faked with inert stand-ins; the
reference files themselves are executed unmodified from where they lie.
"""
import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("IVX_REFERENCE_ROOT", "/root/reference")

_DET = "mmdet3d/models/detectors/imvoxelnet.py"
_NECK = "mmdet3d/models/necks/imvoxelnet.py"


def available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, _DET))


class _Registry:
    def __init__(self):
        self.modules = {}

    def register_module(self, *a, **k):
        def deco(cls):
            self.modules[cls.__name__] = cls
            return cls
        return deco


def _stub_modules():
    import torch

    saved = {}

    def put(name, mod):
        saved[name] = sys.modules.get(name)
        sys.modules[name] = mod

    mmdet = types.ModuleType("mmdet")
    models = types.ModuleType("mmdet.models")
    models.DETECTORS = _Registry()
    models.NECKS = _Registry()
    models.build_backbone = models.build_head = models.build_neck = lambda cfg: None
    dets = types.ModuleType("mmdet.models.detectors")

    class BaseDetector(torch.nn.Module):
        def init_weights(self, pretrained=None):
            pass

    dets.BaseDetector = BaseDetector
    mmdet.models = models
    models.detectors = dets
    mmcv = types.ModuleType("mmcv")
    runner = types.ModuleType("mmcv.runner")

    def auto_fp16(*a, **k):
        def deco(fn):
            return fn
        return deco

    runner.auto_fp16 = auto_fp16
    mmcv.runner = runner
    mmdet3d = types.ModuleType("mmdet3d")
    core = types.ModuleType("mmdet3d.core")
    core.bbox3d2result = lambda *a, **k: None
    mmdet3d.core = core
    for name, mod in (("mmdet", mmdet), ("mmdet.models", models),
                      ("mmdet.models.detectors", dets), ("mmcv", mmcv),
                      ("mmcv.runner", runner), ("mmdet3d", mmdet3d),
                      ("mmdet3d.core", core)):
        put(name, mod)
    return saved


def _restore(saved):
    for name, mod in saved.items():
        if mod is None:
            sys.modules.pop(name, None)
        else:
            sys.modules[name] = mod


_cache = {}


def load():
    """Return ``(detector_module, neck_module)`` of the live reference."""
    if "mods" in _cache:
        return _cache["mods"]
    if not available():
        raise FileNotFoundError(f"reference tree not found at {REFERENCE_ROOT}")
    saved = _stub_modules()
    try:
        out = []
        for i, rel in enumerate((_DET, _NECK)):
            spec = importlib.util.spec_from_file_location(
                f"_ivx_reference_{i}", os.path.join(REFERENCE_ROOT, rel))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            out.append(mod)
    finally:
        _restore(saved)
    _cache["mods"] = tuple(out)
    return _cache["mods"]
