"""ORACLE shim: import-only placeholders for the (missing) pytorch-frame fork."""
class TensorFrame:  # noqa: D101
    pass
class _S(str):
    pass
class stype:  # noqa: D101
    numerical = _S("numerical"); categorical = _S("categorical"); timestamp = _S("timestamp")
    relation = _S("relation"); mask = _S("mask")
from . import nn  # noqa: E402,F401
