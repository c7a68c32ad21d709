"""dml_b200 -- B200-native batch-inference engine behind the reference's models.py API.

The directory name carries a hyphen; import it with
``importlib.import_module("distributed-machine-learning_b200")``.
"""
from . import _lib  # noqa: F401
from ._lib import DmlError, build  # noqa: F401

__all__ = ["DmlError", "build"]
