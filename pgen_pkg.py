"""Import helper: the package directory is ``textsummarization-on-flink_b200`` (hyphenated, after
the reference repository), which ``import`` cannot spell.  ``load()`` registers it as
``textsummarization_on_flink_b200``."""
import importlib.util
import os
import sys

PKG_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "textsummarization-on-flink_b200")
MOD_NAME = "textsummarization_on_flink_b200"


def load():
    if MOD_NAME in sys.modules:
        return sys.modules[MOD_NAME]
    spec = importlib.util.spec_from_file_location(
        MOD_NAME, os.path.join(PKG_DIR, "__init__.py"), submodule_search_locations=[PKG_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[MOD_NAME] = mod
    spec.loader.exec_module(mod)
    return mod
