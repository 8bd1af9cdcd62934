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
_HEADS = ("mmdet3d/models/dense_heads/imvoxel_head.py", "mmdet3d/models/dense_heads/imvoxel_head_v2.py")


def _stub_head_modules(detector_module):
    """Extra stand-ins the two dense-head files import (mmcv 1.2.7 / mmdet 2.10.0 names)."""
    import torch

    mods = {}

    def new(name):
        mods[name] = types.ModuleType(name)
        return mods[name]

    core = new("mmdet.core")

    def multi_apply(func, *args, **kwargs):         # mmdet/core/utils/misc.py: map + transpose
        from functools import partial
        results = map(partial(func, **kwargs) if kwargs else func, *args)
        return tuple(map(list, zip(*results)))

    core.multi_apply = multi_apply
    core.reduce_mean = lambda t: t
    builder = new("mmdet.models.builder")
    builder.HEADS = _Registry()
    builder.build_loss = lambda cfg: None
    cnn = new("mmcv.cnn")

    class Scale(torch.nn.Module):                   # mmcv/cnn/bricks/scale.py: learnable scalar factor
        def __init__(self, scale=1.0):
            super().__init__()
            self.scale = torch.nn.Parameter(torch.tensor(scale, dtype=torch.float))

        def forward(self, x):
            return x * self.scale

    cnn.Scale = Scale
    cnn.bias_init_with_prob = lambda p: 0.0
    cnn.normal_init = lambda *a, **k: None
    det = new("mmdet3d.models.detectors.imvoxelnet")
    det.get_points = detector_module.get_points
    new("mmdet3d.models"), new("mmdet3d.models.detectors"), new("mmdet3d.core.bbox")
    structures = new("mmdet3d.core.bbox.structures")
    structures.rotation_3d_in_axis = None
    post = new("mmdet3d.core.post_processing")
    post.aligned_3d_nms = post.box3d_multiclass_nms = None
    return mods


def load_heads():
    """Return ``(imvoxel_head module, imvoxel_head_v2 module)`` of the live reference."""
    if "heads" in _cache:
        return _cache["heads"]
    detector_module, _ = load()
    saved = _stub_modules()
    extra = _stub_head_modules(detector_module)
    for name, mod in extra.items():
        saved.setdefault(name, sys.modules.get(name))
        sys.modules[name] = mod
    try:
        out = []
        for i, rel in enumerate(_HEADS):
            spec = importlib.util.spec_from_file_location(
                f"_ivx_reference_head_{i}", os.path.join(REFERENCE_ROOT, rel))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            out.append(mod)
    finally:
        _restore(saved)
    _cache["heads"] = tuple(out)
    return _cache["heads"]



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
