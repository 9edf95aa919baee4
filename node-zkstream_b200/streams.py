"""Batch-mode ZKDecodeStream / ZKEncodeStream: the reference's stream interface
(lib/zk-streams.js:25-37,109-119) over the CUDA codec.

Same constructor options ({'fsm': ..., 'isServer': bool}), same packet objects, same
error objects and the same side effect on fsm.zcf_xidMap.  What differs is when work
happens: bytes (or packets) are collected and handed to the GPU in batches.  With the
default batch size of 0 every write() is decoded at once, which reproduces the reference
call for call; a capture replay sets `batchBytes` so that thousands of frames go through
one kernel launch.

The streams are duck-typed like node's: write(), end(), read(), on('data'|'readable'|
'error'|'end', fn).  `fsm` is whatever the reference accepts: an object with
isInState(name) and a zcf_xidMap dict (test/nasty.test.js:298-310).
"""
from . import _abi
from ._lib import Context, XidTable
from .consts import OP_CODES
from .encode import EncodeAssertionError, RequestBatch
from .errors import ZKProtocolError
from .materialise import Materialiser

_default_ctx = None


def default_context():
    """One shared codec context on cuda:0 (created on first use; raises without a GPU)."""
    global _default_ctx
    if _default_ctx is None:
        _default_ctx = Context(0)
    return _default_ctx


def _get(obj, name):
    return obj[name] if isinstance(obj, dict) else getattr(obj, name)


def _in_state(fsm, st):
    f = _get(fsm, "isInState")
    return bool(f(st))


def _xidmap(fsm):
    return _get(fsm, "zcf_xidMap")


class _Emitter:
    def __init__(self):
        self._handlers = {}
        self._queue = []
        self._ended = False

    def on(self, ev, fn):
        self._handlers.setdefault(ev, []).append(fn)
        return self

    def emit(self, ev, *args):
        hs = self._handlers.get(ev, [])
        if ev == "error" and not hs:
            raise args[0]           # node: unhandled 'error' event throws
        for fn in hs:
            fn(*args)

    def push(self, item):
        if self._handlers.get("data"):
            self.emit("data", item)
        else:
            self._queue.append(item)
            self.emit("readable")

    def read(self):
        return self._queue.pop(0) if self._queue else None


def xidtable_from_map(xid_map):
    """fsm.zcf_xidMap ({xid: opcode name}) -> the dense table the decoder takes."""
    m = {}
    for k, v in xid_map.items():
        op = OP_CODES.get(v)
        if op is not None:
            m[int(k)] = op
    return XidTable.from_maps([m])


class ZKDecodeStream(_Emitter):
    """bytes in, packet objects out (lib/zk-streams.js:25-106)."""

    def __init__(self, opts):
        _Emitter.__init__(self)
        if not isinstance(opts, dict):
            raise EncodeAssertionError("options (object) is required")
        if opts.get("fsm") is None:
            raise EncodeAssertionError("options.fsm (object) is required")
        if opts.get("isServer") is not None and not isinstance(opts.get("isServer"), bool):
            raise EncodeAssertionError("options.isServer (bool) is required")
        self.zd_fsm = opts["fsm"]
        self.zd_server = opts.get("isServer") is True
        self.zd_buffer = bytearray()        # bytes not yet split into frames (the reference's zd_buffer)
        self.ctx = opts.get("ctx")
        self.batchBytes = int(opts.get("batchBytes", 0))
        self.handshakeFrames = opts.get("handshakeFrames")   # None: ask the fsm, as the reference does
        self.decodeStrings = opts.get("decodeStrings", True)
        self._stuck = False                 # after BAD_LENGTH the reference never calls cb() again

    @property
    def zd_used(self):
        return len(self.zd_buffer)

    def write(self, chunk, enc=None, cb=None):
        if not isinstance(chunk, (bytes, bytearray, memoryview)):
            raise EncodeAssertionError("chunk (buffer) is required")
        if self._stuck:
            return False
        self.zd_buffer.extend(chunk)
        if len(self.zd_buffer) >= max(self.batchBytes, 1):
            self._decode()
        if cb and not self._stuck:
            cb()
        return True

    _transform = write

    def flush(self):
        if not self._stuck and self.zd_buffer:
            self._decode()

    def end(self, chunk=None):
        if chunk is not None:
            self.write(chunk)
        self.flush()
        self._ended = True
        self.emit("end")

    def _decode(self):
        ctx = self.ctx or default_context()
        data = bytes(self.zd_buffer)
        if self.handshakeFrames is None:
            hs = 0xFFFFFFFF if _in_state(self.zd_fsm, "handshaking") else 0
        else:
            hs = int(self.handshakeFrames)
        flags = [_abi.SESS_SERVER if self.zd_server else 0]
        xt = None if self.zd_server else xidtable_from_map(_xidmap(self.zd_fsm))
        cols = ctx.decode_host(data, [0, len(data)], flags, [hs], xt)
        if self.handshakeFrames is not None:
            self.handshakeFrames = max(0, int(self.handshakeFrames) - cols.n_frames)
        for kind, item in Materialiser(cols, data, self.decodeStrings).events(0):
            if kind == "pkt":
                self.push(item)
            else:
                if item.code == "BAD_LENGTH":
                    self._stuck = True
                self.emit("error", item)
        del self.zd_buffer[: int(cols.sess_consumed[0])]


class ZKEncodeStream(_Emitter):
    """packet objects in, wire bytes out; records xid -> opcode in fsm.zcf_xidMap
    (lib/zk-streams.js:109-148)."""

    def __init__(self, opts):
        _Emitter.__init__(self)
        if not isinstance(opts, dict):
            raise EncodeAssertionError("options (object) is required")
        if opts.get("fsm") is None:
            raise EncodeAssertionError("options.fsm (object) is required")
        if opts.get("isServer") is not None and not isinstance(opts.get("isServer"), bool):
            raise EncodeAssertionError("options.isServer (bool) is required")
        self.ze_fsm = opts["fsm"]
        self.ze_server = opts.get("isServer") is True
        self.ctx = opts.get("ctx")
        self.batchPackets = int(opts.get("batchPackets", 1))
        self._pending = []      # (packet, handshaking at write time)

    def write(self, pkt, enc=None, cb=None):
        if not isinstance(pkt, dict):
            raise EncodeAssertionError("packet (object) is required")
        hs = _in_state(self.ze_fsm, "handshaking")
        if not hs and self.ze_server:
            # the reference calls sub.writeResponse(), which does not exist (zk-streams.js:140)
            raise TypeError("sub.writeResponse is not a function")
        # validate now: the reference throws synchronously from write()
        RequestBatch.from_packets([pkt], hs, self.ze_server)
        self._pending.append((pkt, hs))
        if len(self._pending) >= self.batchPackets:
            self.flush()
        if cb:
            cb()
        return True

    _transform = write

    def flush(self):
        if not self._pending:
            return
        ctx = self.ctx or default_context()
        pkts = [p for p, _ in self._pending]
        hss = [h for _, h in self._pending]
        self._pending = []
        batch = RequestBatch.from_packets(pkts, hss, self.ze_server)
        data, off, status = ctx.encode_host(batch)
        for i, pkt in enumerate(pkts):
            if status[i] != _abi.ENC_OK:
                raise Exception("Unsupported opcode " + str(pkt.get("opcode")))
            self.push(bytes(data[int(off[i]):int(off[i + 1])]))
            if not hss[i]:
                _xidmap(self.ze_fsm)[pkt["xid"]] = pkt.get("opcode")     # zk-streams.js:145

    def end(self, pkt=None):
        if pkt is not None:
            self.write(pkt)
        self.flush()
        self._ended = True
        self.emit("end")
