"""Import helper: the product package lives in `kubernetes-gpu-cluster_b200/`, which is not a
valid Python identifier, so it is loaded by path under the module name `kge_b200`."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG_DIR = os.path.join(ROOT, "kubernetes-gpu-cluster_b200")


def load():
    if "kge_b200" in sys.modules:
        return sys.modules["kge_b200"]
    spec = importlib.util.spec_from_file_location(
        "kge_b200", os.path.join(PKG_DIR, "__init__.py"), submodule_search_locations=[PKG_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["kge_b200"] = mod
    spec.loader.exec_module(mod)
    return mod
