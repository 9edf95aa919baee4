"""TEST INFRASTRUCTURE — ctypes binding of oracle/libzkoracle.so (oracle/zkoracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this.
It reuses the product's ctypes struct mirrors (the oracle emits the same columnar layout
so results compare with array equality) but shares no codec code with the product.
"""
import ctypes as C
import importlib
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
zk = importlib.import_module("node-zkstream_b200")
_abi = zk._abi
Columns = zk.Columns

LIB = os.path.join(HERE, "libzkoracle.so")
_lib = None


def build(force=False):
    src = os.path.join(HERE, "zkoracle.c")
    hdr = os.path.join(ROOT, "include", "zkbatch.h")
    if (force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src)
            or os.path.getmtime(LIB) < os.path.getmtime(hdr)):
        subprocess.run(["gcc", "-O2", "-Wall", "-fPIC", "-shared", "-o", LIB, src, "-lpthread"], check=True)
    return LIB


def load():
    global _lib
    if _lib is None:
        build()
        lib = C.CDLL(LIB)
        P = C.c_void_p
        lib.zko_decode.restype = C.c_int
        lib.zko_decode.argtypes = [P, C.c_uint64, C.POINTER(_abi.Sessions), C.POINTER(_abi.XidTab), C.c_int,
                                   C.c_uint64, C.POINTER(_abi.Decoded), C.POINTER(P)]
        lib.zko_free.argtypes = [P]
        lib.zko_bench_decode.restype = C.c_double
        lib.zko_bench_decode.argtypes = [P, C.c_uint64, C.POINTER(_abi.Sessions), C.POINTER(_abi.XidTab),
                                         C.c_int, C.c_uint64, C.c_int, C.POINTER(C.c_uint64),
                                         C.POINTER(C.c_uint64)]
        lib.zko_digest.restype = C.c_int
        lib.zko_digest.argtypes = [P, C.c_uint64, C.POINTER(_abi.Sessions), C.POINTER(_abi.XidTab), C.c_int,
                                   C.POINTER(C.c_uint64 * 29), C.POINTER(C.c_uint64 * 4)]
        lib.zko_xidtab_build.restype = C.c_int
        lib.zko_xidtab_build.argtypes = [C.POINTER(_abi.Decoded), C.c_uint32, C.POINTER(_abi.XidTab), P]
        lib.zko_session_stats.restype = C.c_int
        lib.zko_session_stats.argtypes = [C.POINTER(_abi.Decoded), C.c_uint32, C.c_uint32,
                                          C.POINTER(_abi.SessionStats)]
        lib.zko_encode.restype = C.c_int
        lib.zko_encode.argtypes = [C.POINTER(_abi.Requests), C.POINTER(P), C.POINTER(C.c_uint64), P, P]
        lib.zko_buf_free.argtypes = [P]
        lib.zko_bench_encode.restype = C.c_double
        lib.zko_bench_encode.argtypes = [C.POINTER(_abi.Requests), C.c_int, C.POINTER(C.c_uint64),
                                         C.POINTER(C.c_uint64)]
        _lib = lib
    return _lib


def _sessions(nbytes, sess_off, flags, hs):
    return zk.Context._sessions(nbytes, sess_off, flags, hs)


def decode(data, sess_off=None, flags=None, handshake_frames=None, xidtab=None, mode=0, chunk=65536):
    """CPU decode of a batch.  mode 0 = in place, 1 = faithful (chunked, copy + memmove per frame)."""
    lib = load()
    buf = np.ascontiguousarray(np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data)
    s, keep = _sessions(buf.size, sess_off, flags, handshake_frames)
    xt = xidtab.struct() if xidtab is not None else None
    out = _abi.Decoded()
    h = C.c_void_p()
    rc = lib.zko_decode(buf.ctypes.data if buf.size else None, buf.size, C.byref(s),
                        C.byref(xt) if xt is not None else None, mode, chunk, C.byref(out), C.byref(h))
    if rc != 0:
        raise ValueError("zko_decode failed: %d" % rc)
    cols = Columns.from_struct(out, s.nsessions, copy=True)
    lib.zko_free(h)
    del keep
    return cols


def bench_decode(data, sess_off, flags, handshake_frames, xidtab, mode, chunk, nthreads):
    """Timed multi-thread CPU decode; returns (seconds, frames)."""
    lib = load()
    buf = np.ascontiguousarray(data, dtype=np.uint8)
    s, keep = _sessions(buf.size, sess_off, flags, handshake_frames)
    xt = xidtab.struct() if xidtab is not None else None
    frames, chk = C.c_uint64(), C.c_uint64()
    sec = lib.zko_bench_decode(buf.ctypes.data, buf.size, C.byref(s), C.byref(xt) if xt is not None else None,
                               mode, chunk, nthreads, C.byref(frames), C.byref(chk))
    del keep
    return sec, frames.value


DIGEST_COLUMNS = ([n for n, _ in _abi.FRAME_COLS + _abi.STAT_COLS + _abi.BUF_COLS + _abi.ACL_COLS] +
                  [n for n, _ in _abi.SESS_COLS])


def digest(data, sess_off, flags, handshake_frames, xidtab, nthreads):
    """Whole-batch decode on `nthreads` threads folded into one order-sensitive 64-bit digest per
    output column (zko_digest).  Returns ({column: digest}, (n_frames, n_stat, n_buf, n_acl))."""
    lib = load()
    buf = np.ascontiguousarray(data, dtype=np.uint8)
    s, keep = _sessions(buf.size, sess_off, flags, handshake_frames)
    xt = xidtab.struct() if xidtab is not None else None
    dig, cnt = (C.c_uint64 * 29)(), (C.c_uint64 * 4)()
    rc = lib.zko_digest(buf.ctypes.data if buf.size else None, buf.size, C.byref(s),
                        C.byref(xt) if xt is not None else None, nthreads, C.byref(dig), C.byref(cnt))
    if rc != 0:
        raise ValueError("zko_digest failed: %d" % rc)
    del keep
    return dict(zip(DIGEST_COLUMNS, [int(x) for x in dig])), tuple(int(x) for x in cnt)


def digest_columns(cols):
    """The same digests computed from a Columns object (numpy), for the tests of zko_digest."""
    M = (1 << 64) - 1

    def mix(v):
        v = v.astype(np.uint64)
        v ^= v >> np.uint64(33); v *= np.uint64(0xff51afd7ed558ccd); v ^= v >> np.uint64(29)
        v *= np.uint64(0xc4ceb9fe1a85ec53); v ^= v >> np.uint64(32)
        return v

    out = {}
    for name in DIGEST_COLUMNS:
        a = getattr(cols, name)
        if a.dtype.itemsize == 4:
            a = a.view(np.uint32)
        else:
            a = a.view(np.uint64)
        w = np.arange(1, a.size + 1, dtype=np.uint64)
        with np.errstate(over="ignore"):
            out[name] = int((mix(a) * w).sum(dtype=np.uint64)) & M
    return out


def xidtab_build(req_cols, layout):
    lib = load()
    d = _abi.Decoded()
    d.n_frames = req_cols.n_frames
    d.meta, d.frame_len = req_cols.meta.ctypes.data, req_cols.frame_len.ctypes.data
    d.session, d.xid = req_cols.session.ctypes.data, req_cols.xid.ctypes.data
    t = layout.struct()
    ops = np.empty(layout.nops, dtype=np.int8)
    lib.zko_xidtab_build(C.byref(d), len(layout.count), C.byref(t), ops.ctypes.data if ops.size else None)
    return zk.XidTable(layout.base, layout.count, ops)


def session_stats(cols, nsessions, flags=0):
    """ZKSession's per-packet folds over decoded response columns -> zk.SessionStats."""
    lib = load()
    d = _abi.Decoded()
    d.n_frames = cols.n_frames
    for name in ("meta", "session", "zxid", "err", "a0", "a1"):
        setattr(d, name, getattr(cols, name).ctypes.data)
    st = zk.SessionStats.empty(nsessions)
    lib.zko_session_stats(C.byref(d), nsessions, flags, C.byref(st.struct()))
    return st


def encode(req):
    """req: RequestBatch.  Returns (bytes ndarray, frame_off, status) like Context.encode_host."""
    lib = load()
    r, keep = req.struct()
    out = C.c_void_p()
    n = C.c_uint64()
    fo = np.zeros(req.n + 1, dtype=np.uint64)
    st = np.zeros(max(req.n, 1), dtype=np.uint8)
    lib.zko_encode(C.byref(r), C.byref(out), C.byref(n), fo.ctypes.data, st.ctypes.data)
    data = np.ctypeslib.as_array(C.cast(out, C.POINTER(C.c_uint8)), shape=(n.value,)).copy() if n.value else np.zeros(0, np.uint8)
    lib.zko_buf_free(out)
    del keep
    return data, fo, st[:req.n]


def bench_encode(req, nthreads):
    """Timed multi-thread CPU encode of a RequestBatch; returns (seconds, frame bytes written)."""
    lib = load()
    r, keep = req.struct()
    nb, chk = C.c_uint64(), C.c_uint64()
    sec = lib.zko_bench_encode(C.byref(r), nthreads, C.byref(nb), C.byref(chk))
    del keep
    return sec, nb.value
