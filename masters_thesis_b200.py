"""Importable alias of the ``masters-thesis_b200`` package directory (a hyphen is not a Python
identifier).  `import masters_thesis_b200[.sub]` resolves to the very same module objects as
importlib.import_module("masters-thesis_b200[.sub]"), so there is one copy of every class."""
import importlib
import importlib.abc
import importlib.util
import sys

_REAL, _ALIAS = "masters-thesis_b200", "masters_thesis_b200"


class _AliasLoader(importlib.abc.Loader):
    def __init__(self, real_name):
        self.real_name = real_name

    def create_module(self, spec):
        return importlib.import_module(self.real_name)

    def exec_module(self, module):
        pass


class _AliasFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        if fullname == _ALIAS or fullname.startswith(_ALIAS + "."):
            real = _REAL + fullname[len(_ALIAS):]
            return importlib.util.spec_from_loader(fullname, _AliasLoader(real))
        return None


if not any(isinstance(f, _AliasFinder) for f in sys.meta_path):
    sys.meta_path.insert(0, _AliasFinder())
_pkg = importlib.import_module(_REAL)
for _name, _mod in list(sys.modules.items()):
    if _name.startswith(_REAL + "."):
        sys.modules[_ALIAS + _name[len(_REAL):]] = _mod
sys.modules[_ALIAS] = _pkg
