"""zkcap: the capture container of include/zkcapture.h — writer, mmap reader, replay driver.

A capture is a list of connections; a connection is a list of segments (request bytes,
response bytes) cut where the connection was quiescent (see the header for why).  The reader
returns numpy views of the mmap'ed file that go straight into Context.replay_host.
"""
import mmap
import os
import struct

import numpy as np

MAGIC = b"ZKCAP\x00\x00\x01"
VERSION = 1
ALIGN = 4096
_HDR = struct.Struct("<8sII" + "Q" * 12 + "16s")     # 128 bytes
assert _HDR.size == 128


class CaptureError(ValueError):
    pass


def _pad(f):
    pos = f.tell()
    rem = (-pos) % ALIGN
    if rem:
        f.write(b"\0" * rem)
    return f.tell()


def write_capture(path, connections, xid_windows=None):
    """connections: iterable of connections, each a list of (req_bytes, rsp_bytes) segments
    (a bare (req, rsp) pair is a connection of one segment).  xid_windows: optional list, one
    entry per segment in order, of (xid_base, xid_count); default (0, frames-in-request-part
    upper bound) is computed by counting length prefixes of the request part."""
    segs = []
    for ci, conn in enumerate(connections):
        if isinstance(conn, tuple) and len(conn) == 2 and isinstance(conn[0], (bytes, bytearray, memoryview, np.ndarray)):
            conn = [conn]
        for k, (rq, rs) in enumerate(conn):
            segs.append((ci, k == 0, bytes(rq), bytes(rs)))
    n = len(segs)
    req_off = np.zeros(n + 1, np.uint64)
    rsp_off = np.zeros(n + 1, np.uint64)
    conn_id = np.zeros(n, np.uint32)
    req_hs = np.zeros(n, np.uint32)
    rsp_hs = np.zeros(n, np.uint32)
    xb = np.zeros(n, np.int32)
    xc = np.zeros(n, np.uint32)
    for i, (ci, first, rq, rs) in enumerate(segs):
        req_off[i + 1] = req_off[i] + len(rq)
        rsp_off[i + 1] = rsp_off[i] + len(rs)
        conn_id[i] = ci
        req_hs[i] = rsp_hs[i] = 1 if first else 0
        if xid_windows is not None:
            xb[i], xc[i] = xid_windows[i]
        else:
            xb[i], xc[i] = 0, _count_frames(rq)
    with open(path, "wb") as f:
        f.write(b"\0" * _HDR.size)
        offs = {}
        for name, arr in (("req_off", req_off), ("rsp_off", rsp_off), ("conn", conn_id), ("xid_base", xb),
                          ("xid_count", xc), ("req_hs", req_hs), ("rsp_hs", rsp_hs)):
            offs[name] = _pad(f)
            f.write(arr.tobytes())
        offs["req"] = _pad(f)
        for _, _, rq, _ in segs:
            f.write(rq)
        offs["rsp"] = _pad(f)
        for _, _, _, rs in segs:
            f.write(rs)
        total = _pad(f)
        f.seek(0)
        f.write(_HDR.pack(MAGIC, VERSION, n, int(req_off[-1]), int(rsp_off[-1]), offs["req_off"], offs["rsp_off"],
                          offs["conn"], offs["xid_base"], offs["xid_count"], offs["req_hs"], offs["rsp_hs"],
                          offs["req"], offs["rsp"], total, b"\0" * 16))
    return n


def _count_frames(b):
    """Upper bound on the xids a request part can carry: its complete frames."""
    n, pos = 0, 0
    while pos + 4 <= len(b):
        ln = int.from_bytes(b[pos:pos + 4], "big", signed=True)
        if ln < 0 or ln > 16 * 1024 * 1024 or pos + 4 + ln > len(b):
            break
        n += 1
        pos += 4 + ln
    return n


class Capture:
    """Read-only view of a zkcap file.  Attributes are numpy views of the mapping: req, rsp
    (uint8), req_off, rsp_off (uint64[n+1]), conn, xid_base, xid_count, req_hs, rsp_hs."""

    def __init__(self, path):
        self._f = open(path, "rb")
        size = os.fstat(self._f.fileno()).st_size
        if size < _HDR.size:
            raise CaptureError("%s: shorter than a zkcap header" % path)
        self._mm = mmap.mmap(self._f.fileno(), 0, access=mmap.ACCESS_READ)
        (magic, version, n, req_bytes, rsp_bytes, o_req_off, o_rsp_off, o_conn, o_xb, o_xc, o_rqhs, o_rshs,
         o_req, o_rsp, total, _) = _HDR.unpack_from(self._mm, 0)
        if magic != MAGIC or version != VERSION:
            raise CaptureError("%s: not a zkcap version %d file" % (path, VERSION))
        if total != size:
            raise CaptureError("%s: %d bytes on disk, header says %d (truncated?)" % (path, size, total))
        self.nsegments = n

        def view(off, count, dt):
            nbytes = count * np.dtype(dt).itemsize
            if off % ALIGN or off + nbytes > size:
                raise CaptureError("%s: section at %d out of bounds or unaligned" % (path, off))
            return np.frombuffer(self._mm, dtype=dt, count=count, offset=off)

        self.req_off = view(o_req_off, n + 1, np.uint64)
        self.rsp_off = view(o_rsp_off, n + 1, np.uint64)
        self.conn = view(o_conn, n, np.uint32)
        self.xid_base = view(o_xb, n, np.int32)
        self.xid_count = view(o_xc, n, np.uint32)
        self.req_hs = view(o_rqhs, n, np.uint32)
        self.rsp_hs = view(o_rshs, n, np.uint32)
        self.req = view(o_req, req_bytes, np.uint8)
        self.rsp = view(o_rsp, rsp_bytes, np.uint8)
        for off, nb, what in ((self.req_off, req_bytes, "request"), (self.rsp_off, rsp_bytes, "response")):
            if n and (off[0] != 0 or off[-1] != nb or (np.diff(off.astype(np.int64)) < 0).any()):
                raise CaptureError("%s: %s segment offsets are not a monotone cover of the section" % (path, what))
        if n > 1 and (np.diff(self.conn.astype(np.int64)) < 0).any():
            raise CaptureError("%s: connection ids are not in order" % path)

    def close(self):
        for name in ("req_off", "rsp_off", "conn", "xid_base", "xid_count", "req_hs", "rsp_hs", "req", "rsp"):
            self.__dict__.pop(name, None)
        self._mm.close()
        self._f.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def replay(self, ctx, on_slice, slice_bytes=0, copy=True):
        """Decode the whole capture through zkb_replay_host; on_slice(first_segment, nsegments,
        req_cols, rsp_cols) as in Context.replay_host (session numbers are segment numbers)."""
        ctx.replay_host(self.req, self.req_off, self.rsp, self.rsp_off, self.xid_base, self.xid_count, on_slice,
                        req_hs=self.req_hs, rsp_hs=self.rsp_hs, slice_bytes=slice_bytes, copy=copy)
