"""Importable alias for the package directory `tensorflow-distributed-on-gke_b200/` (its name is not a valid
Python identifier): `import b200xf` gives that package."""
import importlib.util
import sys
from pathlib import Path

_PKG_DIR = Path(__file__).resolve().parent / "tensorflow-distributed-on-gke_b200"
_spec = importlib.util.spec_from_file_location("b200xf", _PKG_DIR / "__init__.py", submodule_search_locations=[str(_PKG_DIR)])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["b200xf"] = _mod
_spec.loader.exec_module(_mod)
