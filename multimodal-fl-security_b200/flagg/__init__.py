"""flagg -- B200-native robust aggregation engine (Python host side).

``flagg.engine`` binds libflagg.so (include/flagg.h) with ctypes; ``src.defenses`` (next
to this package) mirrors the reference's plugin classes on top of it.
"""
from ._lib import FlaggError, LIB_PATH, MEDIAN, TRIMMED_MEAN, EXPORTS, lib  # noqa: F401
from . import engine  # noqa: F401
from .engine import Comm, Updates  # noqa: F401
