"""String-id registry with the reference's contract (safe_control_gym/utils/registration.py:89-163):
``register(id, entry_point="module:Class", config_entry_point=None)`` and ``make(id, **kwargs)``."""
import importlib

_REGISTRY = {}


def register(id, entry_point, config_entry_point=None, **default_kwargs):
    _REGISTRY[id] = dict(entry_point=entry_point, config_entry_point=config_entry_point, kwargs=default_kwargs)


def _load(name):
    mod, attr = name.split(":")
    return getattr(importlib.import_module(mod), attr)


def make(id, *args, **kwargs):
    if id not in _REGISTRY:
        raise KeyError(f"No registered env with id: {id}")
    spec = _REGISTRY[id]
    merged = dict(spec["kwargs"])
    merged.update(kwargs)
    obj = _load(spec["entry_point"])(*args, **merged)
    obj.spec = spec
    return obj


def get_config(id):
    spec = _REGISTRY[id]
    if spec["config_entry_point"] is None:
        return {}
    return _load(spec["config_entry_point"])


def registered_ids():
    return sorted(_REGISTRY)
