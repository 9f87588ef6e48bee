"""deepsdf_b200 - B200-native (sm_100a) implementation of the DeepSDF auto-decoder hot path.

Python here is host glue only (device memory, streams, autograd wiring, torch.distributed);
all compute is in ``libdsdf_b200.so`` (C ABI in ``include/dsdf.h``).  No CPU path, no fallback.
"""
from ._lib import DsdfError, lib  # noqa: F401
from .spec import NetSpec  # noqa: F401
