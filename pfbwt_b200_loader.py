"""The package directory is called `pfbwt-f_b200` (the reference's name), which is not a Python
identifier; this helper imports its ctypes binding by path."""
import importlib.util
import os

_HERE = os.path.dirname(os.path.abspath(__file__))


def binding():
    spec = importlib.util.spec_from_file_location("pfbwt_f_b200_binding", os.path.join(_HERE, "pfbwt-f_b200", "binding.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod
