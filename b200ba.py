"""Import shim: exposes the package in ``sfm-python_b200/`` (a directory name Python cannot import) as ``b200ba``."""
import importlib.util as _u
import os as _os
import sys as _sys

_dir = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "sfm-python_b200")
_spec = _u.spec_from_file_location("b200ba", _os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = _u.module_from_spec(_spec)
_sys.modules["b200ba"] = _mod
_spec.loader.exec_module(_mod)
