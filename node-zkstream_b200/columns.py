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
        """Bytes of columns written (the output half of the algorithmic traffic)."""
        return sum(getattr(self, n).nbytes for n in self.column_names())
