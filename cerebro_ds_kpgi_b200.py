"""Import alias.  The package directory is `cerebro-ds_kpgi_b200/` (the layout contract's name,
with a hyphen); Python needs an identifier, so `import cerebro_ds_kpgi_b200` lands here and this
module replaces itself with the real package."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cerebro-ds_kpgi_b200")
_spec = importlib.util.spec_from_file_location(
    __name__, os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
