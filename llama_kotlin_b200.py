"""Import alias: the package directory is named `llama.kotlin_b200/` (not a valid
Python identifier), so this module loads it as the package `llama_kotlin_b200`."""
import importlib.util as _ilu
import os as _os
import sys as _sys

_pkg = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "llama.kotlin_b200")
_spec = _ilu.spec_from_file_location("llama_kotlin_b200", _os.path.join(_pkg, "__init__.py"),
                                     submodule_search_locations=[_pkg])
_mod = _ilu.module_from_spec(_spec)
_sys.modules["llama_kotlin_b200"] = _mod
_spec.loader.exec_module(_mod)
