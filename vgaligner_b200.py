"""Import shim: the package directory is `rs-vgaligner_b200/` (hyphenated, as the repo layout
requires), which Python cannot name in an import statement.  `import vgaligner_b200` loads it."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "rs-vgaligner_b200")
_spec = importlib.util.spec_from_file_location("rs_vgaligner_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules.setdefault("rs_vgaligner_b200", _mod)
_spec.loader.exec_module(_mod)
globals().update({k: v for k, v in vars(_mod).items() if not k.startswith("__")})
