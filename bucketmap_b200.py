"""Import shim: the package directory is named ``bucket-map_b200`` (hyphen), so it is loaded here under the
importable name ``bucketmap_b200``.  ``import bucketmap_b200`` from the repo root gives the package."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "bucket-map_b200")
_spec = importlib.util.spec_from_file_location("bucketmap_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["bucketmap_b200"] = _mod
_spec.loader.exec_module(_mod)
