"""Locate the hyphenated package from a bare-name drop-in module."""
import importlib
import os
import sys

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_ROOT = os.path.dirname(_PKG_DIR)
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)


def load(sub):
    return importlib.import_module(os.path.basename(_PKG_DIR) + "." + sub)
