"""memb_b200 — B200-native batched embedding lookup, drop-in for `memb`.

Public surface mirrors python/memb/__init__.py:1-4 of the reference:
Builder, Reader, ReadersUnion, available_compression_strategies.
The native extension `_memb` must be built in-tree (`python build_native.py`);
there is no Python or CPU fallback for the lookup path.
"""
import sys as _sys

try:
    from . import _memb
except ImportError as exc:  # fail loudly: the product is the CUDA path
    raise ImportError(
        "memb_b200._memb is not built. Run `python build_native.py` "
        "(needs nvcc; targets sm_100a only). Original error: %s" % (exc,)
    ) from exc

# the reference package does `import _memb` (python/memb/reader.py:2)
_sys.modules.setdefault("_memb", _memb)

from .builder import Builder  # noqa: E402
from .reader import BaseReader, Reader  # noqa: E402
from .readers_union import ReadersUnion  # noqa: E402

available_compression_strategies = _memb.available_compression_strategies

__all__ = ["Builder", "Reader", "ReadersUnion", "BaseReader", "available_compression_strategies"]
