"""Packets -> struct-of-arrays encode descriptors (zkb_requests, include/zkbatch.h).

Host-side half of the batched ZKEncodeStream: it checks packets the way the reference's
writers do with assert-plus (lib/zk-buffer.js:143-273,384-426 — a bad field is a
synchronous throw) and lays the fields out as columns plus one byte arena.  The bytes of
every frame are produced on the GPU (csrc/zk_encode.cu).
"""
import struct

import numpy as np

from . import _abi
from .consts import CREATE_FLAGS, OP_CODES, PERM_MASKS
from .materialise import LongDate


class EncodeAssertionError(AssertionError):
    """assert-plus failure in the reference's writers (thrown before anything is written)."""
    name = "AssertionError"


def _assert(cond, msg):
    if not cond:
        raise EncodeAssertionError(msg)


def long_to_int64(v, what="long"):
    """writeLong (lib/jute-buffer.js:149-165): Buffer of <= 8 bytes, LongDate, or an integer (the
    reference takes a jsbn BigInteger and uses toByteArray()).  Shorter values are left-padded
    with ZEROS — not sign-extended — exactly as the reference does (quirk 5)."""
    if isinstance(v, LongDate):
        v = v.toBuffer()
    if isinstance(v, (bytes, bytearray, memoryview)):
        b = bytes(v)
    elif isinstance(v, int) and not isinstance(v, bool):
        n = max(1, (v.bit_length() + 8) // 8)       # minimal signed big-endian, as toByteArray()
        b = v.to_bytes(n, "big", signed=True)
    else:
        raise EncodeAssertionError("%s (buffer, LongDate or integer) is required" % what)
    _assert(len(b) <= 8, "%s longer than 8 bytes" % what)
    return struct.unpack(">q", b.rjust(8, b"\x00"))[0]


class RequestBatch:
    """Columns for zkb_encode_*.  Build with from_packets()."""

    def __init__(self):
        self.n = 0
        self.cols = {}
        self.arena = np.zeros(0, np.uint8)
        self.n_elem = 0
        self.n_acl = 0
        self.names = []     # opcode name per packet (None for connect packets)

    @classmethod
    def from_packets(cls, pkts, handshaking=False, is_server=False):
        """pkts: packet dicts as handed to ZKEncodeStream.write().  `handshaking` (bool or one per
        packet) selects writeConnectRequest / writeConnectResponse (lib/zk-streams.js:126-134)."""
        n = len(pkts)
        hs = handshaking if isinstance(handshaking, (list, tuple)) else [bool(handshaking)] * n
        rows = {name: np.zeros(n, dtype=dt) for name, dt in _abi.REQ_COLS}
        arena = bytearray()
        elem, acl = [], []
        names = []

        def put(b):
            off = len(arena)
            arena.extend(b)
            return off, len(b)

        def ustr(v, what):
            _assert(isinstance(v, str), "%s (string) is required" % what)
            return put(v.encode("utf-8"))

        def buf(v, what):
            _assert(isinstance(v, (bytes, bytearray, memoryview)), "%s (buffer) is required" % what)
            return put(bytes(v))

        for i, pkt in enumerate(pkts):
            _assert(isinstance(pkt, dict), "packet (object) is required")
            if hs[i]:
                names.append(None)
                rows["op"][i] = _abi.ENC_CONNECT_RESPONSE if is_server else _abi.ENC_CONNECT_REQUEST
                _assert(isinstance(pkt.get("protocolVersion"), int), "protocolVersion (number) is required")
                _assert(isinstance(pkt.get("timeOut"), int), "timeOut (number) is required")
                rows["a0"][i] = pkt["protocolVersion"]
                rows["a1"][i] = pkt["timeOut"]
                rows["zxid"][i] = long_to_int64(pkt.get("sessionId"), "sessionId")
                if not is_server:
                    rows["zxid2"][i] = long_to_int64(pkt.get("lastZxidSeen"), "lastZxidSeen")
                rows["data_off"][i], rows["data_len"][i] = buf(pkt.get("passwd"), "passwd")
                continue
            _assert(isinstance(pkt.get("xid"), int) and not isinstance(pkt.get("xid"), bool), "xid (number) is required")
            name = pkt.get("opcode")
            _assert(name is None or isinstance(name, str), "opcode (string) is required")
            names.append(name)
            rows["xid"][i] = pkt["xid"]
            op = OP_CODES.get(name)
            if name not in ("GET_CHILDREN", "GET_CHILDREN2", "CREATE", "DELETE", "GET_ACL", "SET_WATCHES",
                            "GET_DATA", "EXISTS", "SET_DATA", "SYNC", "CLOSE_SESSION", "PING"):
                # writeRequest's default branch throws 'Unsupported opcode' (zk-buffer.js:133-134);
                # the device reports it per packet.  Unknown names have no wire integer at all.
                rows["op"][i] = op if op is not None and -128 < op < 128 and op not in (
                    _abi.ENC_CONNECT_REQUEST, _abi.ENC_CONNECT_RESPONSE) else 127
                continue
            rows["op"][i] = op
            if name in ("GET_CHILDREN", "GET_CHILDREN2", "GET_DATA", "EXISTS"):
                rows["path_off"][i], rows["path_len"][i] = ustr(pkt.get("path"), "path")
                _assert(isinstance(pkt.get("watch"), bool), "watch (bool) is required")
                rows["a0"][i] = 1 if pkt["watch"] else 0
            elif name == "CREATE":
                rows["path_off"][i], rows["path_len"][i] = ustr(pkt.get("path"), "path")
                rows["data_off"][i], rows["data_len"][i] = buf(pkt.get("data"), "data")
                acls = pkt.get("acl")
                _assert(isinstance(acls, list) and all(isinstance(x, dict) for x in acls), "acl ([object]) is required")
                rows["vec_first"][i] = len(acl)
                rows["vec_count"][i] = len(acls)
                for line in acls:
                    perms = line.get("perms")
                    _assert(isinstance(perms, list) and all(isinstance(x, str) for x in perms),
                            "permissions ([string]) is required")
                    val = 0
                    for k in perms:                      # writePerms :405-414, any case
                        mask = PERM_MASKS.get(k.upper())
                        _assert(mask is not None, "permission mask for %s (number) is required" % k)
                        val |= mask
                    ident = line.get("id")
                    _assert(isinstance(ident, dict), "id (object) is required")
                    so, sl = ustr(ident.get("scheme"), "id.scheme")
                    io, il = ustr(ident.get("id"), "id")
                    acl.append((val, so, sl, io, il))
                flags = pkt.get("flags")
                _assert(isinstance(flags, list) and all(isinstance(x, str) for x in flags), "flags ([string]) is required")
                fv = 0
                for k in flags:                          # writeCreateRequest :165-172
                    mask = CREATE_FLAGS.get(k)
                    _assert(mask is not None, "unknown flag %s (number) is required" % k)
                    fv |= mask
                rows["a0"][i] = fv
            elif name == "DELETE":
                rows["path_off"][i], rows["path_len"][i] = ustr(pkt.get("path"), "name")
                _assert(isinstance(pkt.get("version"), int), "version (number) is required")
                rows["a0"][i] = pkt["version"]
            elif name in ("GET_ACL", "SYNC"):
                rows["path_off"][i], rows["path_len"][i] = ustr(pkt.get("path"), "path")
            elif name == "SET_DATA":
                rows["path_off"][i], rows["path_len"][i] = ustr(pkt.get("path"), "path")
                rows["data_off"][i], rows["data_len"][i] = buf(pkt.get("data"), "data")
                _assert(isinstance(pkt.get("version"), int), "version (number) is required")
                rows["a0"][i] = pkt["version"]
            elif name == "SET_WATCHES":
                rows["zxid"][i] = long_to_int64(pkt.get("relZxid"), "relZxid")
                ev = pkt.get("events")
                _assert(isinstance(ev, dict), "events (object) is required")
                rows["vec_first"][i] = len(elem)
                counts = []
                for key in ("dataChanged", "createdOrDestroyed", "childrenChanged"):
                    lst = ev.get(key) or []
                    counts.append(len(lst))
                    for j, p in enumerate(lst):
                        elem.append(ustr(p, "%s[%d]" % (key, j)))
                rows["n0"][i], rows["n1"][i] = counts[0], counts[1]
                rows["vec_count"][i] = sum(counts)
        self = cls()
        self.n = n
        self.cols = rows
        self.names = names
        self.arena = np.frombuffer(bytes(arena), dtype=np.uint8) if arena else np.zeros(0, np.uint8)
        self.n_elem, self.n_acl = len(elem), len(acl)
        self.cols["elem_off"] = np.array([e[0] for e in elem], dtype=np.uint64)
        self.cols["elem_len"] = np.array([e[1] for e in elem], dtype=np.int32)
        self.cols["acl_perms"] = np.array([a[0] for a in acl], dtype=np.int32)
        self.cols["acl_scheme_off"] = np.array([a[1] for a in acl], dtype=np.uint64)
        self.cols["acl_scheme_len"] = np.array([a[2] for a in acl], dtype=np.int32)
        self.cols["acl_id_off"] = np.array([a[3] for a in acl], dtype=np.uint64)
        self.cols["acl_id_len"] = np.array([a[4] for a in acl], dtype=np.int32)
        return self

    @classmethod
    def from_columns(cls, cols, data):
        """Decoded client->server frames back into encode descriptors: what a proxy or a capture
        rewriter does between a ZKDecodeStream({isServer: true}) and a ZKEncodeStream.  `cols`
        is the result of decoding `data` with ZKB_SESS_SERVER; every frame the reference would
        have handed on as a packet (OK requests and ConnectRequests) becomes one row, plus the
        header-only PING / CLOSE_SESSION frames, which readRequest rejects (zk-buffer.js:91-92)
        but writeRequest writes (:129-132).  The arena is `data` itself: nothing is copied.
        Returns (batch, frame index per row)."""
        st, kind, op = cols.status, cols.kind, cols.opcode
        req = (st == _abi.ST_OK) & (kind == _abi.KIND_REQUEST)
        con = (st == _abi.ST_OK) & (kind == _abi.KIND_CONNECT_REQUEST)
        hdr_only = ((st == _abi.ST_BAD_DECODE) & (kind == _abi.KIND_REQUEST) & (cols.frame_len == 8) &
                    ((cols.detail & 7) == _abi.CAUSE_UNSUPPORTED) &
                    ((op == OP_CODES["PING"]) | (op == OP_CODES["CLOSE_SESSION"])))
        idx = np.nonzero(req | con | hdr_only)[0]
        n = len(idx)
        rows = {name: np.zeros(n, dtype=dt) for name, dt in _abi.REQ_COLS}
        o = op[idx].astype(np.int8)
        is_con = con[idx]
        rows["op"] = np.where(is_con, np.int8(_abi.ENC_CONNECT_REQUEST), o).astype(np.int8)
        rows["xid"] = np.where(is_con, 0, cols.xid[idx]).astype(np.int32)
        # scalar fields (include/zkbatch.h, FRAMES columns by kind)
        rows["a0"] = np.where(is_con, cols.xid[idx], cols.a0[idx]).astype(np.int32)
        rows["a1"] = np.where(is_con, cols.err[idx], 0).astype(np.int32)
        rows["zxid"] = cols.zxid[idx].astype(np.int64)
        last = (cols.a0[idx].astype(np.int64) << 32) | (cols.a1[idx].astype(np.int64) & 0xFFFFFFFF)
        rows["zxid2"] = np.where(is_con, last, 0).astype(np.int64)
        # buffers: a frame's BUF rows are contiguous and in read order
        b0 = np.searchsorted(cols.buf_frame, idx, "left")
        has_path = np.isin(o, [OP_CODES[k] for k in ("GET_CHILDREN", "GET_CHILDREN2", "GET_DATA", "EXISTS", "CREATE",
                                                      "DELETE", "GET_ACL", "SYNC", "SET_DATA")]) & ~is_con & ~hdr_only[idx]
        has_data = np.isin(o, [OP_CODES["CREATE"], OP_CODES["SET_DATA"]]) & ~is_con
        nb = max(cols.n_buf, 1)
        boff = cols.buf_off if cols.n_buf else np.zeros(1, np.uint64)
        blen = cols.buf_len if cols.n_buf else np.zeros(1, np.int32)
        pr = np.minimum(b0, nb - 1)
        rows["path_off"] = np.where(has_path, boff[pr], 0).astype(np.uint64)
        rows["path_len"] = np.where(has_path, blen[pr], 0).astype(np.int32)
        dr = np.minimum(np.where(is_con, b0, b0 + 1), nb - 1)
        rows["data_off"] = np.where(has_data | is_con, boff[dr], 0).astype(np.uint64)
        rows["data_len"] = np.where(has_data | is_con, blen[dr], 0).astype(np.int32)
        # CREATE: its ACL rows; scheme and id are the BUF rows after path and data
        is_create = (o == OP_CODES["CREATE"]) & ~is_con
        a_first = np.searchsorted(cols.acl_frame, idx, "left")
        rows["vec_first"] = np.where(is_create, a_first, 0).astype(np.uint32)
        rows["vec_count"] = np.where(is_create, np.maximum(cols.a1[idx], 0), 0).astype(np.int32)
        self = cls()
        if cols.n_acl:
            row_of_frame = np.full(cols.n_frames, -1, np.int64)
            row_of_frame[idx] = np.arange(n)
            r = row_of_frame[cols.acl_frame]
            j = np.arange(cols.n_acl) - a_first[r]
            srow = b0[r] + 2 + 2 * j
            self.cols_acl = dict(acl_perms=cols.acl_perms.astype(np.int32),
                                 acl_scheme_off=boff[srow].astype(np.uint64), acl_scheme_len=blen[srow].astype(np.int32),
                                 acl_id_off=boff[srow + 1].astype(np.uint64), acl_id_len=blen[srow + 1].astype(np.int32))
        else:
            self.cols_acl = dict(acl_perms=np.zeros(0, np.int32), acl_scheme_off=np.zeros(0, np.uint64),
                                 acl_scheme_len=np.zeros(0, np.int32), acl_id_off=np.zeros(0, np.uint64),
                                 acl_id_len=np.zeros(0, np.int32))
        # SET_WATCHES: all of the frame's BUF rows are its three path vectors back to back
        is_sw = (o == OP_CODES["SET_WATCHES"]) & ~is_con
        c0 = np.maximum(cols.a0[idx], 0)
        c1 = np.maximum(cols.a1[idx], 0)
        c2 = np.maximum(cols.err[idx], 0)
        tot = np.where(is_sw, c0 + c1 + c2, 0).astype(np.int64)
        e_first = np.cumsum(tot) - tot
        rows["vec_first"] = np.where(is_sw, e_first, rows["vec_first"]).astype(np.uint32)
        rows["vec_count"] = np.where(is_sw, tot, rows["vec_count"]).astype(np.int32)
        rows["n0"] = np.where(is_sw, c0, 0).astype(np.int32)
        rows["n1"] = np.where(is_sw, c1, 0).astype(np.int32)
        if is_sw.any():
            sw_frame = np.zeros(cols.n_frames, bool)
            sw_frame[idx[is_sw]] = True
            m = sw_frame[cols.buf_frame]
            elem_off, elem_len = cols.buf_off[m].astype(np.uint64), cols.buf_len[m].astype(np.int32)
        else:
            elem_off, elem_len = np.zeros(0, np.uint64), np.zeros(0, np.int32)
        self.n = n
        self.cols = rows
        self.cols.update(self.cols_acl)
        del self.cols_acl
        self.cols["elem_off"], self.cols["elem_len"] = elem_off, elem_len
        self.n_elem, self.n_acl = len(elem_off), int(cols.n_acl)
        self.arena = np.ascontiguousarray(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
        self.names = None
        return self, idx

    def struct(self, ptr=lambda a: a.ctypes.data if a.size else None):
        """(zkb_requests struct with host pointers, objects to keep alive)."""
        r = _abi.Requests()
        r.n, r.arena_len, r.n_elem, r.n_acl = self.n, self.arena.size, self.n_elem, self.n_acl
        keep = [self.arena]
        r.arena = ptr(self.arena)
        for name, dt in _abi.REQ_COLS + _abi.ELEM_COLS + _abi.REQ_ACL_COLS:
            a = np.ascontiguousarray(self.cols[name], dtype=dt)
            keep.append(a)
            setattr(r, name, ptr(a))
        return r, keep
