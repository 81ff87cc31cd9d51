"""Importable alias for the package directory `lora-sdr-lightweight-standalone-library-clean_b200/`
(its name contains hyphens, so it cannot be imported with a plain `import` statement)."""
import importlib.util
import os
import sys

_PKG_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)),
                        "lora-sdr-lightweight-standalone-library-clean_b200")
_NAME = "lora_sdr_lightweight_standalone_library_clean_b200"


def _load():
    if _NAME in sys.modules:
        return sys.modules[_NAME]
    spec = importlib.util.spec_from_file_location(_NAME, os.path.join(_PKG_DIR, "__init__.py"),
                                                  submodule_search_locations=[_PKG_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = mod
    spec.loader.exec_module(mod)
    return mod


pkg = _load()
globals().update({k: getattr(pkg, k) for k in dir(pkg) if not k.startswith("__")})
