"""Import shim: `terraform-provider-iterative_b200` is not a valid Python identifier."""
import importlib
import sys

_pkg = importlib.import_module("terraform-provider-iterative_b200")
sys.modules[__name__] = _pkg
