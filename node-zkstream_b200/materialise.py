"""Columns -> the reference's packet objects.

Turns the struct-of-arrays result of the CUDA decoder back into exactly what
ZKDecodeStream pushes (lib/zk-streams.js:81,97): dicts with the reference's keys in the
reference's order, `bytes` where it has Buffers (8-byte longs included, quirk 4),
LongDate for ctime/mtime, opcode/err/type/state as names, and the readPerms
operator-precedence quirk applied to the raw perms (lib/zk-buffer.js:395-403).
"""
import datetime
import struct

import numpy as np

from . import _abi
from .consts import (CREATE_FLAGS, ERR_LOOKUP, NOTIFICATION_TYPE_LOOKUP, OP_CODE_LOOKUP, PERM_MASKS,
                     STATE_LOOKUP)
from .errors import ZKProtocolError

OP = {v: k for k, v in OP_CODE_LOOKUP.items()}


class LongDate:
    """lib/jute-buffer.js:79-95: wraps the 8 wire bytes of a millisecond timestamp."""

    def __init__(self, buffer):
        assert isinstance(buffer, (bytes, bytearray)) and len(buffer) == 8
        self.ld_buffer = bytes(buffer)

    def toDate(self):
        hi, lo = struct.unpack(">II", self.ld_buffer)
        n = lo + 4294967296 * hi                 # unsigned, as :84-85
        return datetime.datetime.fromtimestamp(n / 1000.0, datetime.timezone.utc)

    def toBuffer(self):
        return self.ld_buffer

    def __eq__(self, o):
        return isinstance(o, LongDate) and o.ld_buffer == self.ld_buffer

    def __repr__(self):
        return "LongDate(%s)" % self.ld_buffer.hex()


def long_bytes(v):
    """int64 column value -> the 8 wire bytes readLongBuffer returned."""
    return struct.pack(">q", int(v))


def perms_from_raw(val):
    """readPerms as written: `val & MASK != 0` parses as `val & (MASK != 0)` = `val & 1`,
    so an odd value yields all five names and an even one none (zk-buffer.js:398-401)."""
    return [k for k in PERM_MASKS if (val & int(PERM_MASKS[k] != 0))]


def flags_from_raw(flags):
    """readCreateRequest :152-158"""
    return [k for k, mask in CREATE_FLAGS.items() if (flags & mask) == mask]


def decode_error_text(detail, opcode, handshaking):
    """'Failed to decode ...: ' + e.name + ': ' + e.message (zk-streams.js:74-79,90-95).  The
    RangeError / plain AssertionError wording is that of node 0.12, which the reference targets."""
    cause = detail & 7
    if cause == _abi.CAUSE_RANGE:
        txt = "RangeError: index out of range"
    elif cause == _abi.CAUSE_ASSERT:
        txt = "AssertionError: false == true"
    elif cause == _abi.CAUSE_NO_MATCH:
        txt = "AssertionError: reply packet must match a request"
    elif cause == _abi.CAUSE_UNSUPPORTED:
        txt = "Error: Unsupported opcode " + str(OP_CODE_LOOKUP.get(opcode, "undefined"))
    else:
        txt = "Error: unknown"
    what = "ConnectResponse" if handshaking else "Response"
    return "Failed to decode %s: %s" % (what, txt)


class Materialiser:
    """Walks a Columns result in frame order and yields what the reference stream emits."""

    def __init__(self, cols, data, decode_strings=True):
        self.c = cols
        self.data = data if isinstance(data, (bytes, bytearray, memoryview)) else memoryview(np.ascontiguousarray(data)).cast("B")
        self.decode_strings = decode_strings

    def _bytes(self, row):
        o = int(self.c.buf_off[row])
        return bytes(self.data[o:o + int(self.c.buf_len[row])])

    def _str(self, row):
        b = self._bytes(row)
        if self.decode_strings == "tagged":      # test harnesses: keep ustrings apart from Buffers
            return ("str", b)
        return b.decode("utf-8", "replace") if self.decode_strings else b

    def _stat(self, r):
        c = self.c
        return {
            "czxid": long_bytes(c.czxid[r]), "mzxid": long_bytes(c.mzxid[r]),
            "ctime": LongDate(long_bytes(c.ctime[r])), "mtime": LongDate(long_bytes(c.mtime[r])),
            "version": int(c.version[r]), "cversion": int(c.cversion[r]), "aversion": int(c.aversion[r]),
            "ephemeralOwner": long_bytes(c.ephemeralOwner[r]), "dataLength": int(c.dataLength[r]),
            "numChildren": int(c.numChildren[r]), "pzxid": long_bytes(c.pzxid[r]),
        }

    def _acl(self, n, b, a):
        out = []
        for i in range(n):
            out.append({"perms": perms_from_raw(int(self.c.acl_perms[a + i])),
                        "id": {"scheme": self._str(b + 2 * i), "id": self._str(b + 2 * i + 1)}})
        return out

    def packet(self, f, cur=None):
        """The packet object of FRAMES row f (status must be OK).  cur = (stat, buf, acl) cursors,
        looked up by binary search when not given."""
        c = self.c
        if cur is None:
            cur = (int(np.searchsorted(c.stat_frame, f)), int(np.searchsorted(c.buf_frame, f)),
                   int(np.searchsorted(c.acl_frame, f)))
        s, b, a = cur
        m = int(c.meta[f])
        kind = (m >> 16) & 0xff
        op = ((m >> 8) & 0xff)
        op = op - 256 if op >= 128 else op
        a0, a1 = int(c.a0[f]), int(c.a1[f])
        if kind == _abi.KIND_CONNECT_RESPONSE:
            return {"protocolVersion": int(c.xid[f]), "timeOut": int(c.err[f]),
                    "sessionId": long_bytes(c.zxid[f]), "passwd": self._bytes(b)}
        if kind == _abi.KIND_CONNECT_REQUEST:
            return {"protocolVersion": int(c.xid[f]),
                    "lastZxidSeen": struct.pack(">iI", a0, a1 & 0xffffffff),
                    "timeOut": int(c.err[f]), "sessionId": long_bytes(c.zxid[f]),
                    "passwd": self._bytes(b)}
        name = OP_CODE_LOOKUP.get(op)
        if kind == _abi.KIND_REQUEST:
            pkt = {"xid": int(c.xid[f]), "opcode": name}
            if name in ("GET_CHILDREN", "GET_CHILDREN2", "GET_DATA", "EXISTS"):
                pkt["path"] = self._str(b)
                pkt["watch"] = bool(a0)
            elif name == "CREATE":
                pkt["path"] = self._str(b)
                pkt["data"] = self._bytes(b + 1)
                pkt["acl"] = self._acl(a1, b + 2, a)
                pkt["flags"] = flags_from_raw(a0)
            elif name == "DELETE":
                pkt["path"] = self._str(b)
                pkt["version"] = a0
            elif name in ("GET_ACL", "SYNC"):
                pkt["path"] = self._str(b)
            elif name == "SET_DATA":
                pkt["path"] = self._str(b)
                pkt["data"] = self._bytes(b + 1)
                pkt["version"] = a0
            elif name == "SET_WATCHES":
                n0, n1, n2 = a0, a1, int(c.err[f])
                pkt["relZxid"] = long_bytes(c.zxid[f])
                pkt["events"] = {
                    "dataChanged": [self._str(b + i) for i in range(n0)],
                    "createdOrDestroyed": [self._str(b + n0 + i) for i in range(n1)],
                    "childrenChanged": [self._str(b + n0 + n1 + i) for i in range(n2)],
                }
            return pkt
        # response, zk-buffer.js:284-330
        err = int(c.err[f])
        pkt = {"xid": int(c.xid[f]), "zxid": long_bytes(c.zxid[f]), "err": ERR_LOOKUP.get(err),
               "opcode": name}
        if err != 0:
            return pkt
        if name in ("GET_CHILDREN", "GET_CHILDREN2"):
            pkt["children"] = [self._str(b + i) for i in range(a0)]
            if name == "GET_CHILDREN2":
                pkt["stat"] = self._stat(s)
        elif name == "CREATE":
            pkt["path"] = self._str(b)
        elif name == "GET_ACL":
            pkt["acl"] = self._acl(a0, b, a)
            pkt["stat"] = self._stat(s)
        elif name == "GET_DATA":
            pkt["data"] = self._bytes(b)
            pkt["stat"] = self._stat(s)
        elif name == "NOTIFICATION":
            pkt["type"] = NOTIFICATION_TYPE_LOOKUP.get(a0)
            pkt["state"] = STATE_LOOKUP.get(a1)
            pkt["path"] = self._str(b)
        elif name in ("EXISTS", "SET_DATA"):
            pkt["stat"] = self._stat(s)
        return pkt

    def events(self, session=None):
        """In stream order: ('pkt', dict) for pushed packets, ('error', ZKProtocolError) for
        BAD_DECODE frames and for the BAD_LENGTH that ends a session."""
        c = self.c
        sessions = range(c.nsessions) if session is None else [session]
        for sidx in sessions:
            f0, f1 = int(c.sess_first_frame[sidx]), int(c.sess_first_frame[sidx + 1])
            s = int(np.searchsorted(c.stat_frame, f0))
            b = int(np.searchsorted(c.buf_frame, f0))
            a = int(np.searchsorted(c.acl_frame, f0))
            for f in range(f0, f1):
                m = int(c.meta[f])
                if (m & 0xff) == _abi.ST_OK:
                    yield ("pkt", self.packet(f, (s, b, a)))
                    while s < c.n_stat and c.stat_frame[s] == f:
                        s += 1
                    while b < c.n_buf and c.buf_frame[b] == f:
                        b += 1
                    while a < c.n_acl and c.acl_frame[a] == f:
                        a += 1
                else:
                    op = (m >> 8) & 0xff
                    op = op - 256 if op >= 128 else op
                    kind = (m >> 16) & 0xff
                    hs = kind in (_abi.KIND_CONNECT_REQUEST, _abi.KIND_CONNECT_RESPONSE)
                    yield ("error", ZKProtocolError("BAD_DECODE", decode_error_text(m >> 24, op, hs)))
            if int(c.sess_bad_length_at[sidx]) != _abi.NO_OFFSET:
                yield ("error", ZKProtocolError("BAD_LENGTH", "Invalid ZK packet length"))
