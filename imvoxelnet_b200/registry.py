"""mmdet-compatible registries.

The reference registers its modules in mmdet's ``DETECTORS`` / ``NECKS``
(mmdet3d/models/detectors/imvoxelnet.py:8, necks/imvoxelnet.py:8,70,94,126)
and configs select them by ``type=`` strings.  When mmdet is importable the
drop-in modules register into the real registries (``force=True`` replaces the
reference's classes); otherwise a minimal registry with the same
``register_module`` / ``build`` surface is used so that the ``model`` dicts of
configs/imvoxelnet/*.py still build the neck and detector."""

try:  # pragma: no cover - mmdet is not installed on the build machines
    from mmdet.models import DETECTORS, NECKS
    HAVE_MMDET = True
except Exception:  # ImportError, or mmcv version errors
    HAVE_MMDET = False

    class Registry:
        def __init__(self, name):
            self.name, self.module_dict = name, {}

        def register_module(self, name=None, force=False, module=None):
            def deco(cls):
                key = name or cls.__name__
                if key in self.module_dict and not force:
                    raise KeyError(f"{key} is already registered in {self.name}")
                self.module_dict[key] = cls
                return cls
            return deco(module) if module is not None else deco

        def get(self, key):
            return self.module_dict.get(key)

        def build(self, cfg, **default_args):
            cfg = dict(cfg)
            typ = cfg.pop('type')
            cls = self.get(typ) if isinstance(typ, str) else typ
            if cls is None:
                raise KeyError(f"{typ} is not in the {self.name} registry")
            for k, v in default_args.items():
                cfg.setdefault(k, v)
            return cls(**cfg)

    DETECTORS = Registry('detector')
    NECKS = Registry('neck')


def register(registry):
    """``@register(NECKS)``: works with mmcv's and the local registry."""
    def deco(cls):
        try:
            registry.register_module(force=True)(cls)
        except TypeError:
            registry.register_module()(cls)
        return cls
    return deco


def build_from(registry, cfg):
    if hasattr(registry, 'build'):
        return registry.build(cfg)
    from mmcv.utils import build_from_cfg  # pragma: no cover
    return build_from_cfg(cfg, registry)
