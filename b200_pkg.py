"""Loads the package directory `qvac-ext-lib-llama.cpp_b200/` (its name is not a valid Python identifier) as module `ggml_b200`."""
import importlib.util
import sys
from pathlib import Path

_DIR = Path(__file__).resolve().parent / "qvac-ext-lib-llama.cpp_b200"


def load():
    if "ggml_b200" in sys.modules:
        return sys.modules["ggml_b200"]
    spec = importlib.util.spec_from_file_location("ggml_b200", _DIR / "__init__.py", submodule_search_locations=[str(_DIR)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ggml_b200"] = mod
    spec.loader.exec_module(mod)
    return mod
