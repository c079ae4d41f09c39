"""Import shim: exposes the package in `raytracing-in-a-weekend_b200/` (hyphenated, so not a legal
Python identifier) as the module `rtw_b200`."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "raytracing-in-a-weekend_b200")
_spec = importlib.util.spec_from_file_location("rtw_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["rtw_b200"] = _mod
_spec.loader.exec_module(_mod)

if __name__ == "__main__":
    from rtw_b200.driver import main
    raise SystemExit(main())
