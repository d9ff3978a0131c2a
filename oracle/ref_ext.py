"""Loader for the reference's own extensions built by oracle/build_ref.py (TEST INFRASTRUCTURE).

Returns the pybind module object without registering it under its bare name in
sys.modules, so it cannot shadow this repo's drop-in modules of the same name.
Returns None if the prebuilt .so is absent or cannot be loaded.
"""
import importlib.machinery
import importlib.util
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_cache = {}


def load_ref(name):
    if name in _cache:
        return _cache[name]
    path = os.path.join(_HERE, "_ref", name + ".so")
    mod = None
    if os.path.exists(path):
        try:
            import torch  # noqa: F401  (libtorch must be loaded first)
            loader = importlib.machinery.ExtensionFileLoader(name, path)
            spec = importlib.util.spec_from_loader(name, loader)
            mod = importlib.util.module_from_spec(spec)
            loader.exec_module(mod)
        except Exception as e:  # pragma: no cover
            print(f"[oracle] could not load reference extension {name}: {e}")
            mod = None
    _cache[name] = mod
    return mod
