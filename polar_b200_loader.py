"""Imports the package directory `polar-codes_b200/` (hyphenated, so not importable by name) as `polar_codes_b200`."""
import importlib.util
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))


def load():
    name = "polar_codes_b200"
    if name in sys.modules:
        return sys.modules[name]
    pkg = os.path.join(_ROOT, "polar-codes_b200")
    spec = importlib.util.spec_from_file_location(name, os.path.join(pkg, "__init__.py"), submodule_search_locations=[pkg])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod
