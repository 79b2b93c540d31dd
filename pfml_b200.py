"""Import alias: ``import pfml_b200`` -> the package in the hyphenated directory the build contract names."""
import importlib
import sys

_pkg = importlib.import_module("jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200")
sys.modules[__name__] = _pkg
