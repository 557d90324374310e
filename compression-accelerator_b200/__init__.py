"""libsnappy_b200: B200-native Snappy block compression behind the compress(src, len, dst)
contract of nullromo/compression-accelerator (tests/compression.c:20-27).

The directory name carries a hyphen (it follows the reference repository's name), so load the
package with ``__graft_entry__.load_package()`` (importlib, module name
``compression_accelerator_b200``) rather than a plain ``import``.
"""
from .api import (BLOCK, EXACT, FAST, Context, SnappyB200Error, compress, default_context, lib, debug_stats,  # noqa: F401
                  max_compressed_length, rocc_compress, rocc_uncompress, varint32, LIB_PATH, SYMBOLS, RoccDevice)
