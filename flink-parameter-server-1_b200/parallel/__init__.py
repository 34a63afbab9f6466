from .engine import HostEngine
from .streams import DataStream, as_stream

__all__ = ["HostEngine", "DataStream", "as_stream"]
