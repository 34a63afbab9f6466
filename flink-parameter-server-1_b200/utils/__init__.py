from .eof import EOFHandler, WrapWithEOF, flat_map_with_eof
from .sleep_blocker import block

__all__ = ["EOFHandler", "WrapWithEOF", "flat_map_with_eof", "block"]
