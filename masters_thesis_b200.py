"""Importable alias of the ``masters-thesis_b200`` package directory (a hyphen is not a Python identifier)."""
import importlib
import sys

_pkg = importlib.import_module("masters-thesis_b200")
sys.modules[__name__] = _pkg
