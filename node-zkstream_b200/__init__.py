"""zkstream_b200 — B200-native batch codec for ZooKeeper jute frames.

The data-parallel hot path of node-zkstream (frame splitting + packet decode, and the
symmetric request encode) as hand-written sm_100a CUDA behind a plain C ABI
(include/zkbatch.h, libzkbatch.so), plus a host-side mirror of the reference's
ZKDecodeStream / ZKEncodeStream / ZKBuffer interface.  There is no CPU codec in this
package: without the built library and a CUDA device every call raises.
"""
from . import _abi, consts
from ._lib import Context, XidTable, ZKBatchError, load, LIB_PATH
from .columns import Columns
from .errors import ZKProtocolError, ZKError
from .materialise import LongDate, Materialiser
from .encode import RequestBatch
from .streams import ZKDecodeStream, ZKEncodeStream
from .zkbuffer import JuteBuffer, ZKBuffer
from .session import SessionStats, SessionState

__all__ = ["Context", "XidTable", "ZKBatchError", "Columns", "load", "LIB_PATH", "consts",
           "ZKProtocolError", "ZKError", "LongDate", "Materialiser", "RequestBatch",
           "ZKDecodeStream", "ZKEncodeStream", "JuteBuffer", "ZKBuffer",
           "SessionStats", "SessionState"]
