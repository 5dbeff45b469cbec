"""Import shim: ``import blf_b200`` loads the package that lives in the directory
``bipedal-locomotion-framework_b200/`` (not importable by name because of the hyphens)."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "bipedal-locomotion-framework_b200")
_spec = importlib.util.spec_from_file_location("blf_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["blf_b200"] = _mod
_spec.loader.exec_module(_mod)
