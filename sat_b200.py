"""Importable alias for the package directory ``self-attention-tacotron_b200/``.

The directory name is fixed by the repo layout contract and is not a valid
Python identifier, so this module loads it with importlib and re-exports it as
``sat_b200`` (``import sat_b200; sat_b200.modules.SelfAttention`` ...).
"""
import importlib.util
import os
import sys

_PKG_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "self-attention-tacotron_b200")
_spec = importlib.util.spec_from_file_location(
    "sat_b200", os.path.join(_PKG_DIR, "__init__.py"), submodule_search_locations=[_PKG_DIR])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["sat_b200"] = _mod
_spec.loader.exec_module(_mod)
