"""``ADVERSARIES`` registry and ``build_adversary`` -- the API of the reference's
``models/builder.py`` (/root/reference/models/builder.py:6-11), so that configs such as
``adversary_cfg=dict(type='VoxelPerturber', sensor_error_bound=0.2, ...)``
(/root/reference/configs/adversarial/adversarial-second_hv_secfpn_8xb6-80e_kitti-3d-3class.py:77-82)
build the B200 perturber unchanged.

When mmengine is importable its ``Registry`` is used (scope ``'models'`` as in the reference, so
``custom_imports`` picks the class up); otherwise a minimal stand-in with the same ``register_module`` /
``build`` surface is used -- mmengine is not installed in the build image.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, Optional

try:  # pragma: no cover - mmengine is absent from the build image
    from mmengine.registry import Registry  # type: ignore
except Exception:  # noqa: BLE001

    class Registry:  # type: ignore[no-redef]
        """The part of ``mmengine.registry.Registry`` the reference uses."""

        def __init__(self, name: str, parent: Optional["Registry"] = None, scope: Optional[str] = None):
            self.name, self.scope, self.parent = name, scope, parent
            self.module_dict: Dict[str, Any] = {}

        def register_module(self, name: Optional[str] = None, force: bool = False,
                            module: Optional[type] = None) -> Callable:
            def deco(cls):
                key = name or cls.__name__
                if key in self.module_dict and not force:
                    raise KeyError(f"{key} is already registered in {self.name}")
                self.module_dict[key] = cls
                return cls

            return deco(module) if module is not None else deco

        def get(self, key: str):
            return self.module_dict.get(key)

        def build(self, cfg: Dict[str, Any]):
            if not isinstance(cfg, dict) or "type" not in cfg:
                raise KeyError("cfg must be a dict with a 'type' key")
            args = dict(cfg)
            typ = args.pop("type")
            cls = self.module_dict.get(typ) if isinstance(typ, str) else typ
            if cls is None:
                raise KeyError(f"{typ} is not in the {self.name} registry")
            return cls(**args)


ADVERSARIES = Registry("adversaries", parent=None, scope="models")


def build_adversary(cfg: Dict[str, Any]):
    """Build an adversary from its config dict (reference: models/builder.py:8-11)."""
    return ADVERSARIES.build(cfg)
