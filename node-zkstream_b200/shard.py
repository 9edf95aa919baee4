"""Multi-GPU use: sessions are independent (the xid map is per connection,
lib/connection-fsm.js:39, and the splitter state per stream, lib/zk-streams.js:29-31), so a
batch is cut into contiguous runs of sessions, one per rank, and each rank decodes its run
on its own GPU with no collective.  merge() puts per-rank results back into one table when
a caller wants that; nothing on the data path needs it."""
import numpy as np

from . import _abi
from .columns import Columns


def partition_sessions(sess_off, world):
    """[(s0, s1)] * world: contiguous session ranges with near-equal byte counts."""
    off = np.asarray(sess_off, dtype=np.uint64)
    n = len(off) - 1
    total = int(off[-1])
    cuts = [0]
    for r in range(1, world):
        target = total * r // world
        s = int(np.searchsorted(off, target, side="left"))
        s = min(max(s, cuts[-1]), n)
        cuts.append(s)
    cuts.append(n)
    return [(cuts[i], cuts[i + 1]) for i in range(world)]


def shard_inputs(data, sess_off, flags, hs, xidtab, s0, s1):
    """The slice of a batch that holds sessions [s0, s1): (bytes, offsets, flags, hs, xid table)."""
    from ._lib import XidTable
    off = np.asarray(sess_off, dtype=np.uint64)
    lo, hi = int(off[s0]), int(off[s1])
    xt = None
    if xidtab is not None:
        o0 = int(xidtab.off[s0]) if s0 < len(xidtab.off) else xidtab.nops
        o1 = int(xidtab.off[s1 - 1]) + int(xidtab.count[s1 - 1]) if s1 > s0 else o0
        xt = XidTable(xidtab.base[s0:s1], xidtab.count[s0:s1], xidtab.ops[o0:o1])
    return (data[lo:hi], off[s0:s1 + 1] - off[s0],
            None if flags is None else np.asarray(flags)[s0:s1],
            None if hs is None else np.asarray(hs)[s0:s1], xt)


def merge(parts, byte_bases, session_bases):
    """Concatenate per-shard Columns into one, shifting byte offsets, frame and session indices."""
    out = Columns()
    out.nsessions = sum(p.nsessions for p in parts)
    fbase = np.cumsum([0] + [p.n_frames for p in parts])
    for cnt, cols in (("n_frames", _abi.FRAME_COLS), ("n_stat", _abi.STAT_COLS), ("n_buf", _abi.BUF_COLS),
                      ("n_acl", _abi.ACL_COLS)):
        setattr(out, cnt, sum(getattr(p, cnt) for p in parts))
        for name, dt in cols:
            pieces = []
            for i, p in enumerate(parts):
                a = getattr(p, name)
                if name in ("frame_off", "buf_off"):
                    a = a + np.uint64(byte_bases[i])
                elif name == "session":
                    a = a + np.uint32(session_bases[i])
                elif name in ("stat_frame", "buf_frame", "acl_frame"):
                    a = a + np.uint32(fbase[i])
                pieces.append(a)
            setattr(out, name, np.concatenate(pieces) if pieces else np.zeros(0, dt))
    out.sess_consumed = np.concatenate([p.sess_consumed + np.uint64(byte_bases[i]) for i, p in enumerate(parts)])
    bad = []
    for i, p in enumerate(parts):
        b = p.sess_bad_length_at.copy()
        b[b != _abi.NO_OFFSET] += np.uint64(byte_bases[i])
        bad.append(b)
    out.sess_bad_length_at = np.concatenate(bad)
    first = [p.sess_first_frame[:-1] + np.uint32(fbase[i]) for i, p in enumerate(parts)]
    out.sess_first_frame = np.concatenate(first + [np.array([fbase[-1]], dtype=np.uint32)])
    return out
