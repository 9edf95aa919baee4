"""Loads libzkbatch.so and wraps its C ABI (include/zkbatch.h) for Python.

The library is the only codec: if it is missing or no CUDA device is usable every
entry point here raises; nothing falls back to a CPU implementation.
"""
import ctypes as C
import os

import numpy as np

from . import _abi
from .columns import Columns

_HERE = os.path.dirname(os.path.abspath(__file__))
# ZKBATCH_LIB points at an alternative build of the same library (e.g. a profiling build)
LIB_PATH = os.environ.get("ZKBATCH_LIB") or os.path.join(_HERE, "libzkbatch.so")
_lib = None


class ZKBatchError(RuntimeError):
    def __init__(self, code, msg):
        RuntimeError.__init__(self, "zkbatch error %d: %s" % (code, msg))
        self.code = code


def load():
    """dlopen the in-tree libzkbatch.so and bind every symbol the header declares."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s is missing: run `python node-zkstream_b200/build.py` "
                          "(there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, restype, argtypes in _abi.SYMBOLS:
        fn = getattr(lib, name)   # AttributeError if the symbol is not exported
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def _ptr(a):
    return a.ctypes.data if a is not None and a.size else None


def _arr(a, dt, n=None):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dtype=dt)
    if n is not None and a.shape != (n,):
        raise ValueError("expected %d entries, got %r" % (n, a.shape))
    return a


class XidTable:
    """Batched fsm.zcf_xidMap: per session a dense run of xids -> wire opcode (or OP_NONE)."""

    def __init__(self, base, count, ops=None):
        self.base = np.ascontiguousarray(base, dtype=np.int32)
        self.count = np.ascontiguousarray(count, dtype=np.uint32)
        self.off = np.zeros(len(self.count), dtype=np.uint64)
        if len(self.count):
            self.off[1:] = np.cumsum(self.count[:-1], dtype=np.uint64)
        self.nops = int(self.count.sum())
        if ops is None:
            ops = np.full(self.nops, _abi.OP_NONE, dtype=np.int8)
        self.ops = np.ascontiguousarray(ops, dtype=np.int8)
        assert self.ops.shape == (self.nops,)

    @classmethod
    def from_maps(cls, maps):
        """maps: per session {xid: wire opcode}; only non-negative dense-ish ranges are kept."""
        base, count, chunks = [], [], []
        for m in maps:
            ks = [k for k in m if k >= 0]
            if not ks:
                base.append(0); count.append(0); continue
            lo, hi = min(ks), max(ks)
            row = np.full(hi - lo + 1, _abi.OP_NONE, dtype=np.int8)
            for k in ks:
                row[k - lo] = m[k]
            base.append(lo); count.append(len(row)); chunks.append(row)
        ops = np.concatenate(chunks) if chunks else np.zeros(0, dtype=np.int8)
        return cls(base, count, ops)

    def struct(self):
        t = _abi.XidTab()
        t.base, t.count, t.off, t.ops = _ptr(self.base), _ptr(self.count), _ptr(self.off), _ptr(self.ops)
        t.nops = self.nops
        return t


class Context:
    """A codec context bound to one CUDA device (zkb_ctx)."""

    def __init__(self, device=0):
        self._lib = load()
        h = C.c_void_p()
        rc = self._lib.zkb_ctx_create(int(device), C.byref(h))
        if rc != 0:
            raise ZKBatchError(rc, "zkb_ctx_create(device=%d) failed: no usable CUDA device "
                                   "(this codec has no CPU path)" % device)
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            self._lib.zkb_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise ZKBatchError(rc, self._lib.zkb_last_error(self._h).decode("utf-8", "replace"))

    @property
    def kernel_launches(self):
        return int(self._lib.zkb_kernel_launches(self._h))

    @property
    def split_fixups(self):
        return int(self._lib.zkb_split_fixups(self._h))

    @property
    def row_mispredicts(self):
        """Decode passes repeated with look-back placement (see zkb_row_mispredicts)."""
        return int(self._lib.zkb_row_mispredicts(self._h))

    @property
    def last_slow_frames(self):
        return int(self._lib.zkb_last_slow_frames(self._h))

    @property
    def stream(self):
        return self._lib.zkb_ctx_stream(self._h)

    def set_timing(self, on=True):
        self._check(self._lib.zkb_ctx_set_timing(self._h, 1 if on else 0))

    def set_replay_compact(self, on=True):
        """zkb_replay_host leaves the derivable columns out of the download (Columns.restore_derived
        puts them back): see zkb_ctx_set_replay_compact."""
        self._check(self._lib.zkb_ctx_set_replay_compact(self._h, 1 if on else 0))

    def last_phase_ms(self):
        """(split, emit, decode) device milliseconds of the last decode call (timing must be on)."""
        out = (C.c_float * 3)()
        self._check(self._lib.zkb_last_phase_ms(self._h, C.byref(out)))
        return tuple(float(x) for x in out)

    def sync(self):
        self._check(self._lib.zkb_ctx_sync(self._h))

    # ---- decode -----------------------------------------------------------------
    @staticmethod
    def _sessions(nbytes, sess_off, flags, handshake_frames):
        if sess_off is None:
            sess_off = [0, nbytes]
        off = _arr(sess_off, np.uint64)
        n = len(off) - 1
        fl = _arr(flags, np.uint8, n)
        hs = _arr(handshake_frames, np.uint32, n)
        s = _abi.Sessions()
        s.nsessions = n
        s.off, s.flags, s.handshake_frames = _ptr(off), _ptr(fl), _ptr(hs)
        if n == 0:
            s.off = off.ctypes.data
        return s, (off, fl, hs)

    def decode_host(self, data, sess_off=None, flags=None, handshake_frames=None, xidtab=None, copy=True):
        """Decode a batch held in host memory (bytes / numpy uint8).  Returns Columns."""
        buf = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
        buf = np.ascontiguousarray(buf, dtype=np.uint8)
        s, keep = self._sessions(buf.size, sess_off, flags, handshake_frames)
        xt = xidtab.struct() if xidtab is not None else None
        out = _abi.Decoded()
        rc = self._lib.zkb_decode_host(self._h, _ptr(buf), buf.size, C.byref(s),
                                       C.byref(xt) if xt is not None else None, C.byref(out))
        self._check(rc)
        del keep
        return Columns.from_struct(out, s.nsessions, copy=copy)

    def replay_host(self, req, req_off, rsp, rsp_off, xid_base, xid_count, on_slice, req_flags=None,
                    rsp_flags=None, req_hs=None, rsp_hs=None, slice_bytes=0, copy=True):
        """Pipelined paired replay (zkb_replay_host).  on_slice(first_session, nsessions, req_cols,
        rsp_cols) is called once per slice, in order; with copy=False the Columns are views that
        die when on_slice returns.  Request streams default to ZKB_SESS_SERVER."""
        req = np.ascontiguousarray(req, dtype=np.uint8)
        rsp = np.ascontiguousarray(rsp, dtype=np.uint8)
        n = len(req_off) - 1
        if req_flags is None:
            req_flags = np.full(n, _abi.SESS_SERVER, dtype=np.uint8)
        s1, k1 = self._sessions(req.size, req_off, req_flags, req_hs)
        s2, k2 = self._sessions(rsp.size, rsp_off, rsp_flags, rsp_hs)
        xb = np.ascontiguousarray(xid_base, dtype=np.int32)
        xc = np.ascontiguousarray(xid_count, dtype=np.uint32)
        err = []

        def _cb(user, first, ns, rq, rs):
            try:
                on_slice(int(first), int(ns), Columns.from_struct(rq.contents, int(ns), copy=copy),
                         Columns.from_struct(rs.contents, int(ns), copy=copy))
                return 0
            except Exception as e:      # never let an exception cross the C frame
                err.append(e)
                return -1

        cb = _abi.SLICE_CB(_cb)
        rc = self._lib.zkb_replay_host(self._h, _ptr(req), req.size, C.byref(s1), _ptr(rsp), rsp.size,
                                       C.byref(s2), xb.ctypes.data, xc.ctypes.data, int(slice_bytes), cb, None)
        del k1, k2
        if err:
            raise err[0]
        self._check(rc)

    def decode_dev(self, d_bytes, nbytes, d_sess, d_xid, d_out):
        """Raw device-pointer call (structs from _abi, filled by the caller)."""
        rc = self._lib.zkb_decode_dev(self._h, d_bytes, nbytes, C.byref(d_sess),
                                      C.byref(d_xid) if d_xid is not None else None, C.byref(d_out))
        return rc

    def xidtab_build_host(self, req_cols, layout):
        """xid table for the response streams paired with the decoded request streams `req_cols`."""
        d = _abi.Decoded()
        d.n_frames = req_cols.n_frames
        d.meta, d.frame_len = _ptr(req_cols.meta), _ptr(req_cols.frame_len)
        d.session, d.xid = _ptr(req_cols.session), _ptr(req_cols.xid)
        t = layout.struct()
        ops = np.empty(layout.nops, dtype=np.int8)
        self._check(self._lib.zkb_xidtab_build_host(self._h, C.byref(d), len(layout.count), C.byref(t), _ptr(ops)))
        return XidTable(layout.base, layout.count, ops)

    # ---- session-level reductions ------------------------------------------------
    def session_stats_host(self, cols, nsessions=None, flags=0):
        """ZKSession's per-packet folds (max zxid, notification counts) over decoded response
        columns `cols` -> session.SessionStats."""
        from .session import SessionStats
        n = cols.nsessions if nsessions is None else nsessions
        d = _abi.Decoded()
        d.n_frames = cols.n_frames
        for name in ("meta", "session", "zxid", "a0", "a1"):
            setattr(d, name, _ptr(getattr(cols, name)))
        st = SessionStats.empty(n)
        cs = st.struct()
        self._check(self._lib.zkb_session_stats_host(self._h, C.byref(d), n, flags, C.byref(cs)))
        return st

    # ---- encode -----------------------------------------------------------------
    def encode_host(self, req):
        """req: encode.RequestBatch.  Returns (bytes ndarray, frame_off ndarray, status ndarray)."""
        r, keep = req.struct()
        out, fo, st = C.c_void_p(), C.c_void_p(), C.c_void_p()
        n = C.c_uint64()
        self._check(self._lib.zkb_encode_host(self._h, C.byref(r), C.byref(out), C.byref(fo), C.byref(st),
                                              C.byref(n)))
        del keep
        from .columns import _view
        return (_view(out.value, n.value, "u1").copy(), _view(fo.value, req.n + 1, "u8").copy(),
                _view(st.value, req.n, "u1").copy())
