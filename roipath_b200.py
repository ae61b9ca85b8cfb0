"""Importable alias of the package directory `instancesegmentation-jittor_b200/` (its name
contains a hyphen, so it cannot be imported with a plain `import` statement).

    import roipath_b200 as rp      # with the repo root on sys.path
"""
import importlib.util
import os
import sys

_PKG_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "instancesegmentation-jittor_b200")
_NAME = "instancesegmentation_jittor_b200"


def _load():
    if _NAME in sys.modules:
        return sys.modules[_NAME]
    spec = importlib.util.spec_from_file_location(
        _NAME, os.path.join(_PKG_DIR, "__init__.py"), submodule_search_locations=[_PKG_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = mod
    spec.loader.exec_module(mod)
    return mod


_pkg = _load()
globals().update({k: getattr(_pkg, k) for k in _pkg.__all__})
package = _pkg
