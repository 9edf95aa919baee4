"""Columnar (struct-of-arrays) decode result: numpy views over a zkb_decoded struct.

Layout and meaning of every column: include/zkbatch.h.
"""
import ctypes as C

import numpy as np

from . import _abi

_TABLES = (("n_frames", _abi.FRAME_COLS), ("n_stat", _abi.STAT_COLS), ("n_buf", _abi.BUF_COLS),
           ("n_acl", _abi.ACL_COLS))


def _view(ptr, n, dt):
    dt = np.dtype(dt)
    if n == 0 or not ptr:
        return np.zeros(0, dtype=dt)
    buf = (C.c_uint8 * (n * dt.itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dt, count=n)


class Columns:
    """All tables of one decode call.  Attributes are numpy arrays named as in zkb_decoded."""

    def __init__(self):
        self.nsessions = 0
        self.n_frames = self.n_stat = self.n_buf = self.n_acl = 0

    @classmethod
    def from_struct(cls, d, nsessions, copy=True):
        self = cls()
        self.nsessions = nsessions
        for cnt, cols in _TABLES:
            n = int(getattr(d, cnt))
            setattr(self, cnt, n)
            for name, dt in cols:
                v = _view(getattr(d, name), n, dt)
                setattr(self, name, v.copy() if copy else v)
        for name, dt in _abi.SESS_COLS:
            n = nsessions + 1 if name == "sess_first_frame" else nsessions
            v = _view(getattr(d, name), n, dt)
            setattr(self, name, v.copy() if copy else v)
        if getattr(d, "buf_rel", None):          # compact replay result
            v = _view(d.buf_rel, self.n_buf, "u4")
            self.buf_rel = v.copy() if copy else v
        return self

    def restore_derived(self, sess_byte_off, first_session=0):
        """Put back the columns a compact replay result leaves out (zkb_ctx_set_replay_compact):
        frame_off, session, stat_frame / buf_frame / acl_frame and buf_off.  `sess_byte_off` = the byte
        offsets of this slice's sessions in the caller's buffer ([nsessions + 1] or [nsessions]),
        `first_session` = the slice's first session (the full result numbers sessions globally)."""
        nf, ns = self.n_frames, self.nsessions
        first = self.sess_first_frame.astype(np.int64)
        per = np.diff(first)
        local = np.repeat(np.arange(ns, dtype=np.int64), per)
        self.session = (local + first_session).astype(np.uint32)
        # frames of a session lie back to back from the session's first byte
        lens = self.frame_len.astype(np.int64) + 4
        excl = np.cumsum(lens) - lens
        base = np.zeros(ns, dtype=np.int64)
        has = per > 0
        base[has] = excl[first[:-1][has]]
        start = np.asarray(sess_byte_off, dtype=np.uint64)[:ns].astype(np.int64)
        self.frame_off = (start[local] + excl - base[local]).astype(np.uint64)
        # rows per frame, from what the frame is (include/zkbatch.h, "decoded output")
        ok = (self.meta & 0xff) == 0
        op = ((self.meta >> 8) & 0xff).astype(np.uint8).view(np.int8).astype(np.int64)
        kind = (self.meta >> 16) & 0xff
        a0 = np.maximum(self.a0.astype(np.int64), 0)
        a1 = np.maximum(self.a1.astype(np.int64), 0)
        er = self.err.astype(np.int64)
        nb = np.zeros(nf, np.int64); nst = np.zeros(nf, np.int64); na = np.zeros(nf, np.int64)
        rsp = ok & (kind == _abi.KIND_RESPONSE) & (er == 0)
        nb[rsp & np.isin(op, (8, 12))] = a0[rsp & np.isin(op, (8, 12))]
        m = rsp & (op == 6)
        nb[m] = 2 * a0[m]; na[m] = a0[m]
        nb[rsp & np.isin(op, (1, 4, 0))] = 1
        nst[rsp & np.isin(op, (12, 6, 4, 3, 5))] = 1
        req = ok & (kind == _abi.KIND_REQUEST)
        nb[req & np.isin(op, (8, 12, 4, 3, 2, 6, 9))] = 1
        nb[req & (op == 5)] = 2
        m = req & (op == 1)
        nb[m] = 2 + 2 * a1[m]; na[m] = a1[m]
        m = req & (op == 101)
        nb[m] = a0[m] + a1[m] + np.maximum(er[m], 0)
        nb[ok & (kind >= _abi.KIND_CONNECT_RESPONSE)] = 1
        if int(nb.sum()) != self.n_buf or int(nst.sum()) != self.n_stat or int(na.sum()) != self.n_acl:
            raise ValueError("restore_derived: row counts do not add up (%d/%d buf, %d/%d stat, %d/%d acl)" % (
                nb.sum(), self.n_buf, nst.sum(), self.n_stat, na.sum(), self.n_acl))
        frames = np.arange(nf, dtype=np.uint32)
        self.buf_frame = np.repeat(frames, nb)
        self.stat_frame = np.repeat(frames, nst)
        self.acl_frame = np.repeat(frames, na)
        self.buf_off = self.frame_off[self.buf_frame] + self.buf_rel.astype(np.uint64)
        return self

    # meta word accessors
    @property
    def status(self):
        return (self.meta & 0xff).astype(np.uint8)

    @property
    def opcode(self):
        return ((self.meta >> 8) & 0xff).astype(np.uint8).view(np.int8)

    @property
    def kind(self):
        return ((self.meta >> 16) & 0xff).astype(np.uint8)

    @property
    def detail(self):
        return ((self.meta >> 24) & 0xff).astype(np.uint8)

    def column_names(self):
        names = [n for _, cols in _TABLES for n, _ in cols]
        return names + [n for n, _ in _abi.SESS_COLS]

    def diff(self, other):
        """Names of columns (or counts) that differ from `other`; empty list = bit-exact."""
        bad = []
        for cnt, _ in _TABLES:
            if getattr(self, cnt) != getattr(other, cnt):
                bad.append("%s: %d != %d" % (cnt, getattr(self, cnt), getattr(other, cnt)))
        if bad:
            return bad
        for name in self.column_names():
            a, b = getattr(self, name), getattr(other, name)
            if a.shape != b.shape or a.dtype != b.dtype or not np.array_equal(a, b):
                where = ""
                if a.shape == b.shape and a.size:
                    idx = np.flatnonzero(a != b)
                    where = " first at row %d: %r != %r" % (idx[0], a[idx[0]], b[idx[0]])
                bad.append(name + where)
        return bad

    def nbytes_columns(self):
        """Bytes of columns delivered (the output half of the algorithmic traffic)."""
        extra = self.buf_rel.nbytes if hasattr(self, "buf_rel") else 0
        return sum(getattr(self, n).nbytes for n in self.column_names()) + extra
