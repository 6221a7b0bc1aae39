"""Import shim: the package directory is named ``vlut.cpp_b200`` (not a valid Python identifier), so this
module loads it under the importable name ``vlut_b200`` and replaces itself with it in ``sys.modules``."""
import importlib.util
import sys
from pathlib import Path

_pkg_dir = Path(__file__).resolve().parent / "vlut.cpp_b200"
_spec = importlib.util.spec_from_file_location("vlut_b200", _pkg_dir / "__init__.py",
                                               submodule_search_locations=[str(_pkg_dir)])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["vlut_b200"] = _mod
_spec.loader.exec_module(_mod)
