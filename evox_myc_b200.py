"""Import shim: makes the directory ``evox-myc_b200/`` importable as ``evox_myc_b200``.

The package directory carries the reference's repository name, which contains a hyphen and so
cannot be written in an import statement; this module loads it under a legal name.
"""
import importlib.util as _ilu
import os as _os
import sys as _sys

_dir = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "evox-myc_b200")
_spec = _ilu.spec_from_file_location(__name__, _os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = _ilu.module_from_spec(_spec)
_sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
