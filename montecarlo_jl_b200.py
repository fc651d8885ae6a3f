"""Import shim: the package directory is `montecarlo.jl_b200/` (the name the project
layout fixes); a dot cannot appear in a Python module name, so this module loads that
directory as the package `montecarlo_jl_b200`."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "montecarlo.jl_b200")
_spec = importlib.util.spec_from_file_location(
    "montecarlo_jl_b200", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["montecarlo_jl_b200"] = _mod
_spec.loader.exec_module(_mod)
