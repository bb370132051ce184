# -*- coding: utf-8 -*-
"""Import shim: the package directory is `acoss-1_b200/` (not a valid identifier), so this
module loads it under the importable name `acoss_b200` (submodules included)."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "acoss-1_b200")
_spec = importlib.util.spec_from_file_location("acoss_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["acoss_b200"] = _mod
_spec.loader.exec_module(_mod)
