"""Imports the package directory `qwen3-tts-jetson_b200/` (not a valid identifier) as the module
`qwen3_tts_jetson_b200`."""
import importlib.util
import os
import sys

_NAME = "qwen3_tts_jetson_b200"
_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "qwen3-tts-jetson_b200")


def load():
    if _NAME in sys.modules:
        return sys.modules[_NAME]
    spec = importlib.util.spec_from_file_location(_NAME, os.path.join(_DIR, "__init__.py"),
                                                  submodule_search_locations=[_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = mod
    spec.loader.exec_module(mod)
    return mod
