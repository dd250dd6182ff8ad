"""String-id registry with the reference's contract (safe_control_gym/utils/registration.py:11-163):
``register(id, entry_point="module:Class" | callable, config_entry_point=None | dict | "pkg:file.yaml" | "module:dict_name")``,
``make(id, *args, **kwargs)`` (the instance learns its ``spec``), ``get_config(id)``, ``spec(id)``."""
import copy
import importlib
import os


def _load(name):
    mod, attr = name.split(":")
    return getattr(importlib.import_module(mod), attr)


class Spec:
    """registration.py:32-86."""

    def __init__(self, id, entry_point=None, config_entry_point=None, **default_kwargs):
        self.id, self.entry_point, self.config_entry_point, self.kwargs = id, entry_point, config_entry_point, default_kwargs

    def __getitem__(self, key):   # (round-1 callers read the spec like a dict)
        return getattr(self, key)

    def get_config(self):
        """registration.py:44-66: a dict as is; "pkg:file.yaml" loaded from the package directory; "module:dict_name" imported."""
        cep = self.config_entry_point
        if cep is None:
            return {}
        if isinstance(cep, dict):
            return cep
        if isinstance(cep, str):
            if cep.endswith(".yaml"):
                import yaml
                mod_name, fname = cep.split(":")
                mod = importlib.import_module(mod_name)
                with open(os.path.join(os.path.dirname(os.path.abspath(mod.__file__)), fname)) as f:
                    return yaml.safe_load(f)
            return _load(cep)
        raise Exception("Config type {} is not supported.".format(cep))

    def make(self, *args, **kwargs):
        if self.entry_point is None:
            raise Exception("Attempting to make deprecated env {}.".format(self.id))
        merged = dict(self.kwargs)
        merged.update(kwargs)
        cls = self.entry_point if callable(self.entry_point) else _load(self.entry_point)
        obj = cls(*args, **merged)
        spec = copy.deepcopy(self)
        if hasattr(obj, "spec"):
            obj.instance_spec = spec
        else:
            obj.spec = spec
        return obj


_REGISTRY = {}


def register(id, **kwargs):
    if id in _REGISTRY:
        raise Exception("Cannot re-register id: {}".format(id))
    _REGISTRY[id] = Spec(id, **kwargs)


def spec(id):
    try:
        return _REGISTRY[id]
    except KeyError:
        raise Exception("Key not found in registry.")


def make(id, *args, **kwargs):
    if id not in _REGISTRY:
        raise KeyError(f"No registered env with id: {id}")
    return _REGISTRY[id].make(*args, **kwargs)


def get_config(id):
    return spec(id).get_config()


def registered_ids():
    return sorted(_REGISTRY)
