import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def nsb():
    """The product package (its directory name has a hyphen, hence importlib)."""
    return importlib.import_module("navierstokes-parallel_b200")
