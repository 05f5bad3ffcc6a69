"""Model registry with the semantics of `mon`'s `ModelFactory`
(src/mon/core/factory.py:233-436; global `MODELS`, src/mon/globals.py:1224).

`MODELS[arch][name] = cls`; `MODELS.build(name=..., arch=..., config={...})` tries
the raw / kebab / snake / Pascal spellings of `name` across all archs, exactly as
the reference does.  When `mon` is importable, `register_into_mon()` installs the
B200 classes over the reference's entries (`replace=True`), so
`mon.MODELS.build(config={"name": "zero_dce++_re", ...})` returns the drop-in.
"""
from __future__ import annotations

import copy
import inspect
import re
from typing import Any

__all__ = ["ModelFactory", "MODELS", "register_into_mon", "kebabize", "snakeize", "pascalize"]


def _words(s: str) -> list[str]:
    s = re.sub(r"([a-z0-9])([A-Z])", r"\1 \2", s)
    s = re.sub(r"([A-Z]+)([A-Z][a-z])", r"\1 \2", s)
    return [w for w in re.split(r"[\s_\-]+", s) if w]


def kebabize(s: str | None) -> str | None:
    return None if s is None else "-".join(w.lower() for w in _words(s))


def snakeize(s: str | None) -> str | None:
    return None if s is None else "_".join(w.lower() for w in _words(s))


def pascalize(s: str | None) -> str | None:
    return None if s is None else "".join(w[:1].upper() + w[1:] for w in _words(s))


class ModelFactory(dict):
    """`dict` of `arch -> {name -> class}` (factory.py:233-248)."""

    def __init__(self, name: str = "Models"):
        super().__init__()
        self.factory_name = name

    @property
    def archs(self) -> list[str]:
        return list(self.keys())

    @property
    def models(self) -> list[str]:
        return [k for d in self.values() if isinstance(d, dict) for k in d]

    def register(self, name: str = None, arch: str = None, module: Any = None, replace: bool = False):
        """Decorator or plain call (factory.py:257-298)."""
        if not (name is None or isinstance(name, str)):
            raise TypeError(f"`name` must be a `str`, but got {type(name)}.")
        if module:
            self.register_module(module, name, arch, replace)
            return module

        def _register(cls):
            self.register_module(cls, name, arch, replace)
            return cls
        return _register

    def register_module(self, module_cls: Any, module_name: str = None, arch_name: str = None,
                        replace: bool = False):
        """A second registration under the same (arch, name) is ignored unless `replace`
        (factory.py:300-328)."""
        if not inspect.isclass(module_cls):
            raise ValueError(f"`module_cls` must be a class interface, but got {type(module_cls)}.")
        module_name = module_name or kebabize(module_cls.__name__)
        arch_name = arch_name or kebabize(getattr(module_cls, "arch", None)) or kebabize(module_cls.__name__)
        self.setdefault(arch_name, {})
        if replace or module_name not in self[arch_name]:
            self[arch_name][module_name] = module_cls

    def build(self, name: str = None, arch: str = None, config: dict = None, to_dict: bool = False, **kwargs):
        """factory.py:330-387."""
        if (name is None and config is None) or (name is None and config and "name" not in config):
            return None
        if config:
            config_ = copy.deepcopy(config)
            name = name or config_.pop("name", None)
            kwargs |= config_
        arch = arch or name
        found = False
        if name:
            for n in [name, kebabize(name), snakeize(name), pascalize(name)]:
                for a, models in self.items():
                    if n in models:
                        name, arch, found = n, a, True
                        break
                if found:
                    break
        if not found or arch not in self or name not in self[arch]:
            raise ValueError(f"`arch` and `name` must be a valid keyword inside the registry, "
                             f"but got {arch} and {name}.")
        obj = self[arch][name](**kwargs)
        if getattr(obj, "name", None) is None:
            obj.name = snakeize(name)
        return {f"{name}": obj} if to_dict else obj

    def build_instances(self, configs: list, to_dict: bool = False, **kwargs):
        if configs is None:
            return None
        if not isinstance(configs, list):
            raise ValueError(f"`configs` must be a `list`, but got {type(configs)}.")
        objs = {} if to_dict else []
        for config in copy.deepcopy(configs):
            if isinstance(config, str):
                obj = self.build(name=config, to_dict=to_dict, **kwargs)
            elif isinstance(config, dict):
                name = config.pop("name", None)
                arch = config.pop("arch", None)
                obj = self.build(name=name, arch=arch, to_dict=to_dict, **(kwargs | config))
            else:
                raise ValueError(f"each item of `configs` must be a `str` or `dict`, got {type(config)}.")
            if obj is not None:
                objs.update(obj) if to_dict else objs.append(obj)
        return objs or None


MODELS = ModelFactory("Models")


def register_into_mon() -> bool:
    """Install the drop-ins over `mon.MODELS["zero_dce"][...]` when `mon` imports.  Returns
    whether it did.  (`import mon` needs lightning/plum/... which may be absent.)"""
    try:
        import mon  # type: ignore
    except Exception:
        return False
    for arch, models in MODELS.items():
        for name, cls in models.items():
            mon.MODELS.register(name=name, arch=arch, module=cls, replace=True)
    return True
