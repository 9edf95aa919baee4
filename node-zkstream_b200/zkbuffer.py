"""JuteBuffer / ZKBuffer with the reference's method names (lib/jute-buffer.js,
lib/zk-buffer.js).  The primitives (readInt, writeBuffer, ...) are plain cursor
arithmetic on the host; every PACKET-level reader and writer goes through the CUDA codec
as a batch of one frame, so there is a single implementation of the schemas: the kernels."""
import struct

from . import _abi
from .consts import OP_CODES
from .encode import EncodeAssertionError, RequestBatch, long_to_int64
from .materialise import LongDate, Materialiser, decode_error_text
from .streams import default_context, xidtable_from_map


class RangeError(Exception):
    """node's Buffer RangeError (out-of-bounds readInt32BE)."""
    name = "RangeError"


class JuteBuffer:
    def __init__(self, opts):
        if not isinstance(opts, dict):
            raise EncodeAssertionError("options (object) is required")
        buf = opts.get("buffer")
        if buf is not None and not isinstance(buf, (bytes, bytearray, memoryview)):
            raise EncodeAssertionError("options.buffer (buffer) is required")
        self.jb_buffer = bytearray(buf) if buf is not None else bytearray(1024)
        self.jb_size = len(self.jb_buffer)
        self.jb_offset = 0
        self.ctx = opts.get("ctx")

    # ---- lib/jute-buffer.js:23-44
    def toBuffer(self):
        return bytes(self.jb_buffer[: self.jb_offset])

    def atEnd(self):
        return self.jb_offset >= len(self.jb_buffer)

    def remainder(self):
        return bytes(self.jb_buffer[self.jb_offset:])

    def skip(self, n):
        self.jb_offset += n

    def expand(self):
        self.jb_size *= 2
        self.jb_buffer.extend(bytes(self.jb_size - len(self.jb_buffer)))

    # ---- read primitives :46-109
    def readByte(self):
        if self.jb_offset + 1 > len(self.jb_buffer):
            raise RangeError("index out of range")
        v = struct.unpack_from(">b", self.jb_buffer, self.jb_offset)[0]
        self.jb_offset += 1
        return v

    def readBool(self):
        assert self.jb_offset < self.jb_size
        v = self.jb_buffer[self.jb_offset]
        self.jb_offset += 1
        assert v in (0, 1)
        return v == 1

    def readInt(self):
        if self.jb_offset + 4 > len(self.jb_buffer):
            raise RangeError("index out of range")
        v = struct.unpack_from(">i", self.jb_buffer, self.jb_offset)[0]
        self.jb_offset += 4
        return v

    def readLongBuffer(self):
        assert self.jb_offset + 8 <= self.jb_size
        v = bytes(self.jb_buffer[self.jb_offset:self.jb_offset + 8])
        self.jb_offset += 8
        return v

    def readLongDate(self):
        return LongDate(self.readLongBuffer())

    def readBuffer(self):
        n = self.readInt()
        if n < 0:
            n = 0
        assert self.jb_offset + n <= self.jb_size
        v = bytes(self.jb_buffer[self.jb_offset:self.jb_offset + n])
        self.jb_offset += n
        return v

    def readUString(self):
        return self.readBuffer().decode("utf-8", "replace")

    # ---- write primitives :111-165,181-189
    def _room(self, n):
        while self.jb_offset + n > self.jb_size:
            self.expand()

    def writeBool(self, v, name=None):
        if not isinstance(v, bool):
            raise EncodeAssertionError("%s (bool) is required" % (name or "value"))
        self.writeByte(1 if v else 0)

    def writeByte(self, v, name=None):
        self._room(1)
        struct.pack_into(">b", self.jb_buffer, self.jb_offset, v)
        self.jb_offset += 1

    def writeBuffer(self, v, name=None):
        if not isinstance(v, (bytes, bytearray, memoryview)):
            raise EncodeAssertionError("%s (buffer) is required" % (name or "value"))
        self._room(len(v) + 4)
        if len(v) == 0:
            self.writeInt(-1)           # :127-130
            return
        self.writeInt(len(v))
        self.jb_buffer[self.jb_offset:self.jb_offset + len(v)] = bytes(v)
        self.jb_offset += len(v)

    def writeUString(self, v, name=None):
        if not isinstance(v, str):
            raise EncodeAssertionError("%s (string) is required" % (name or "value"))
        self.writeBuffer(v.encode("utf-8"))

    def writeInt(self, v, name=None):
        if not isinstance(v, int) or isinstance(v, bool):
            raise EncodeAssertionError("%s (number) is required" % (name or "value"))
        self._room(4)
        struct.pack_into(">i", self.jb_buffer, self.jb_offset, v)
        self.jb_offset += 4

    def writeLong(self, v, name=None):
        self._room(8)
        struct.pack_into(">q", self.jb_buffer, self.jb_offset, long_to_int64(v))
        self.jb_offset += 8

    def writeLengthPrefixed(self, cb):
        at = self.jb_offset
        self._room(4)
        self.jb_offset += 4
        ret = cb(self)
        struct.pack_into(">I", self.jb_buffer, at, self.jb_offset - at - 4)
        return ret


class ZKBuffer(JuteBuffer):
    """lib/zk-buffer.js:17-20 — packet readers and writers."""

    # ---- readers: one frame through zkb_decode_host ------------------------------------
    def _read(self, server, handshaking, xid_map=None):
        ctx = self.ctx or default_context()
        payload = bytes(self.jb_buffer[self.jb_offset:self.jb_size])
        data = struct.pack(">I", len(payload)) + payload
        xt = xidtable_from_map(xid_map) if xid_map is not None else None
        cols = ctx.decode_host(data, [0, len(data)], [_abi.SESS_SERVER if server else 0],
                               [1 if handshaking else 0], xt)
        assert cols.n_frames == 1
        m = int(cols.meta[0])
        if (m & 0xff) != _abi.ST_OK:
            op = (m >> 8) & 0xff
            op = op - 256 if op >= 128 else op
            text = decode_error_text(m >> 24, op, handshaking).split(": ", 1)[1]
            kind, msg = text.split(": ", 1)
            if kind == "RangeError":
                raise RangeError(msg)
            if kind == "AssertionError":
                raise AssertionError(msg)
            raise Exception(msg)
        mat = Materialiser(cols, data)
        pkt = mat.packet(0, (0, 0, 0))
        self.jb_offset += self._consumed(cols)
        return pkt

    @staticmethod
    def _consumed(cols):
        """Bytes of the payload the reader walked over (where jb_offset ends up)."""
        m = int(cols.meta[0])
        kind = (m >> 16) & 0xff
        op = (m >> 8) & 0xff
        op = op - 256 if op >= 128 else op
        start = 4
        end = {_abi.KIND_RESPONSE: 16, _abi.KIND_REQUEST: 8,
               _abi.KIND_CONNECT_RESPONSE: 16, _abi.KIND_CONNECT_REQUEST: 24}[kind]
        if kind == _abi.KIND_REQUEST and op == OP_CODES["SET_WATCHES"]:
            end = 8 + 8 + 12
        if kind == _abi.KIND_RESPONSE and int(cols.err[0]) == 0:
            if op in (OP_CODES["GET_CHILDREN"], OP_CODES["GET_CHILDREN2"], OP_CODES["GET_ACL"]):
                end = 20
            if op == OP_CODES["NOTIFICATION"]:
                end = 24
        if kind == _abi.KIND_REQUEST and op == OP_CODES["CREATE"]:
            end = 8
        if cols.n_buf:
            end = max(end, int(cols.buf_off[-1]) + int(cols.buf_len[-1]) - start)
        if kind == _abi.KIND_REQUEST:
            if op in (OP_CODES["GET_CHILDREN"], OP_CODES["GET_CHILDREN2"], OP_CODES["GET_DATA"], OP_CODES["EXISTS"]):
                end += 1
            elif op in (OP_CODES["DELETE"], OP_CODES["SET_DATA"]):
                end += 4
            elif op == OP_CODES["CREATE"]:
                if cols.n_acl == 0:
                    end += 4        # the ACL count
                end += 4            # flags
        if cols.n_stat:
            end += 68
        return end

    def readConnectRequest(self):
        return self._read(True, True)

    def readConnectResponse(self):
        return self._read(False, True)

    def readRequest(self):
        return self._read(True, False)

    def readResponse(self, xidMap):
        if not isinstance(xidMap, dict):
            raise EncodeAssertionError("xidMap (object) is required")
        return self._read(False, False, xidMap)

    # ---- writers: one packet through zkb_encode_host -----------------------------------
    def _write(self, pkt, handshaking, server):
        ctx = self.ctx or default_context()
        batch = RequestBatch.from_packets([pkt], handshaking, server)
        data, off, status = ctx.encode_host(batch)
        if status[0] != _abi.ENC_OK:
            # the reference has written xid and opcode before it throws (zk-buffer.js:98-99,134)
            self.writeInt(pkt["xid"])
            op = OP_CODES.get(pkt.get("opcode"))
            if op is None:
                raise EncodeAssertionError("value (number) is required")
            self.writeInt(op)
            raise Exception("Unsupported opcode " + str(pkt.get("opcode")))
        body = bytes(data[4:int(off[1])])
        self._room(len(body))
        self.jb_buffer[self.jb_offset:self.jb_offset + len(body)] = body
        self.jb_offset += len(body)

    def writeConnectRequest(self, pkt):
        self._write(pkt, True, False)

    def writeConnectResponse(self, pkt):
        self._write(pkt, True, True)

    def writeRequest(self, pkt):
        self._write(pkt, False, False)
