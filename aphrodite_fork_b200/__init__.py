"""Import shim: the package directory is `aphrodite-fork_b200/` (not a Python identifier);
this module makes it importable as `aphrodite_fork_b200`."""
import os as _os

__path__.insert(0, _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "aphrodite-fork_b200"))

from ._loader import load, library_paths, c_api  # noqa: E402,F401
from . import _custom_ops as ops  # noqa: E402,F401
from .paged_attn import PagedAttention, PagedAttentionMetadata  # noqa: E402,F401
from . import tp  # noqa: E402,F401
from . import marlin_utils  # noqa: E402,F401
from .tp import TPGroup, shard_heads, shard_marlin_weight  # noqa: E402,F401
from .decode import DecodeConfig, DecodeModel  # noqa: E402,F401
