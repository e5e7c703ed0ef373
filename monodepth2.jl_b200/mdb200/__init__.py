"""mdb200 -- B200-native view-synthesis loss path for Monodepth2.jl (host-side mirror).

Importing this package loads libmdb200.so (built in-tree); it raises ImportError if the
library is missing -- there is no CPU fallback.
"""
from . import _lib  # noqa: F401  (loads the shared library, fails loudly if absent)
from .api import *  # noqa: F401,F403
from .api import _apply_mask  # noqa: F401
