"""The five synthetic workloads of BASELINE.json (`configs`, made concrete in SURVEY.md §8d) at
benchmark size, as plain arrays ready for the C ABI: one or two *directions* (a batch of session
byte streams each) plus what the caller has to say about them.

    cfg1  1 M GET_DATA replies (256-byte payload + Stat, 348-byte frames), one session
    cfg2  mixed request/response replay with xid->opcode pairing, ~10 GB         (the headline)
    cfg3  tiny frames: notifications (36-44 B) and ping replies (20 B), one giant session
    cfg4  large znodes: 1 MB GET_DATA payloads and GET_CHILDREN2 replies with 10^4-10^5 children
    cfg5  the cfg2 mix at 10^5 frames per connection (5 x 10^4 each way), 1250 connections per GPU

bench.py times them; the generators are csrc/zkgen.c (multi-threaded C) and synth.py.
"""
import numpy as np

from . import _abi, synth
from ._lib import XidTable

NAMES = ("cfg1", "cfg2", "cfg3", "cfg4", "cfg5")
LABEL = {
    "cfg1": "cfg1 1M getData replies (256 B payload + Stat)",
    "cfg2": "cfg2 mixed request/response replay, xid->opcode pairing",
    "cfg3": "cfg3 tiny frames (notifications 36-44 B, ping replies 20 B), one session",
    "cfg4": "cfg4 large znodes (1 MB payloads, 10^4-10^5-child vectors)",
    "cfg5": "cfg5 sharded replay, 1250 connections x 10^5 frames per GPU (cfg2 mix)",
}
# wire bytes of one connection (both directions) per frame pair of the cfg2 mix, measured
CFG2_BYTES_PER_FRAME_PAIR = 7.79e6 / 20000
DEFAULT_GB = {"cfg1": 0.348, "cfg2": 10.0, "cfg3": 1.13, "cfg4": 8.0, "cfg5": 1250 * 50000 * CFG2_BYTES_PER_FRAME_PAIR / 1e9}
FRAMES_PER_SESSION = {"cfg2": 20000, "cfg5": 50000}


class Direction:
    """One batch of session streams that all run the same way."""

    def __init__(self, name, data, off, flags=None, hs=None, xid=None, frames_hint=None):
        self.name = name
        self.data = data                      # uint8 [nbytes]
        self.off = np.ascontiguousarray(off, dtype=np.uint64)
        self.flags = flags                    # uint8 [S] or None (client side)
        self.hs = hs
        self.xid = xid                        # None | "paired" (built from the request direction) | XidTable
        self.frames_hint = frames_hint        # (frames, stat, buf, acl) capacity guess
        self.nsessions = len(self.off) - 1
        self.nbytes = int(self.off[-1])


class Workload:
    def __init__(self, cfg, directions, layout=None, truth_ops=None, info=None):
        self.cfg = cfg
        self.label = LABEL[cfg]
        self.directions = directions
        self.layout = layout                  # paired: XidTable layout (base/count) of the connections
        self.truth_ops = truth_ops            # paired: the generator's own xid -> opcode table
        self.info = info or {}
        self.nbytes = sum(d.nbytes for d in directions)
        self.nsessions = directions[0].nsessions

    @property
    def paired(self):
        return self.layout is not None

    def truth_table(self):
        return XidTable(self.layout.base, self.layout.count, self.truth_ops)


def _pinned_alloc(pinned, keep):
    if not pinned:
        return lambda n: np.empty(n, dtype=np.uint8)
    import torch

    def alloc(n):
        t = torch.empty(n + 64, dtype=torch.uint8, pin_memory=True)
        keep.append(t)
        return t.numpy()[:n]
    return alloc


def _paired(cfg, seed, nsessions, frames, pinned):
    keep = []
    pc = synth.native_cfg2(seed, nsessions, frames, alloc=_pinned_alloc(pinned, keep))
    nf = nsessions * (frames + 1)
    req = Direction("request streams", pc.req, pc.req_off, pc.req_flags, None, None,
                    (nf, 16, 2 * nf, nf // 4 + 16))
    rsp = Direction("response streams", pc.rsp, pc.rsp_off, pc.rsp_flags, None, "paired", (nf, nf, 4 * nf, 16))
    layout = XidTable(np.zeros(nsessions, np.int32), np.full(nsessions, frames, np.uint32))
    w = Workload(cfg, [req, rsp], layout, pc.ops, {"sessions": nsessions, "frames_per_session": frames})
    w._keep = keep
    return w


def _into(alloc, a):
    out = alloc(a.size)
    out[:] = a
    return out


def make(cfg, seed, gb=None, pinned=False, nsessions=None):
    """The workload of `cfg` at about `gb` GB of wire bytes (None = the size BASELINE names).
    `nsessions` (cfg1/cfg3 only) cuts the same number of frames into that many sessions — used for
    the CPU baseline's sample, which shards sessions over threads."""
    gb = DEFAULT_GB[cfg] if gb is None else gb
    keep = []
    alloc = _pinned_alloc(pinned, keep)
    if cfg in ("cfg2", "cfg5"):
        frames = FRAMES_PER_SESSION[cfg]
        ns = max(1, int(round(gb * 1e9 / (CFG2_BYTES_PER_FRAME_PAIR * frames))))
        return _paired(cfg, seed, ns, frames, pinned)
    if cfg == "cfg1":
        total = max(1000, int(gb * 1e9 / 348))
        ns = nsessions or 1
        per = total // ns
        parts = [synth.native_cfg1(seed + i, per, 256) for i in range(ns)]
        off = np.zeros(ns + 1, np.uint64)
        off[1:] = np.cumsum([p.size for p in parts])
        data = alloc(int(off[-1]))
        for i, p in enumerate(parts):
            data[int(off[i]):int(off[i + 1])] = p
        xt = XidTable(np.zeros(ns, np.int32), np.full(ns, per, np.uint32), np.full(ns * per, 4, np.int8))
        d = Direction("getData replies", data, off, None, None, xt, (ns * (per + 1), ns * per + 16, ns * (per + 1) + 16, 16))
        w = Workload(cfg, [d], info={"sessions": ns, "frames_per_session": per + 1})
        w._keep = keep
        return w
    if cfg == "cfg3":
        total = max(1000, int(gb * 1e9 / 35.3))
        ns = nsessions or 1
        per = total // ns
        parts = [synth.native_cfg3(seed + i, per) for i in range(ns)]
        off = np.zeros(ns + 1, np.uint64)
        off[1:] = np.cumsum([p.size for p in parts])
        data = alloc(int(off[-1]))
        for i, p in enumerate(parts):
            data[int(off[i]):int(off[i + 1])] = p
        nf = ns * (per + 1)
        d = Direction("notification / ping streams", data, off, None, None, None, (nf, 16, nf, 16))
        w = Workload(cfg, [d], info={"sessions": ns, "frames_per_session": per + 1})
        w._keep = keep
        return w
    if cfg == "cfg4":
        c = synth.cfg4_large(40, seed=seed % 100000)
        one = np.frombuffer(c.data, dtype=np.uint8)
        reps = max(1, int(gb * 1e9 // one.size) + (1 if gb * 1e9 > one.size else 0))
        data = alloc(one.size * reps)
        for r in range(reps):
            data[r * one.size:(r + 1) * one.size] = one
        off = np.arange(reps + 1, dtype=np.uint64) * np.uint64(one.size)
        xt = XidTable.from_maps([c.xidmaps[0]] * reps)
        nkids = one.size // 12
        d = Direction("large-znode replies", data, off, None, None, xt, (reps * 48, reps * 48, reps * nkids, 16))
        w = Workload(cfg, [d], info={"sessions": reps, "frames_per_session": 41,
                                     "note": "one 40-frame session (20 x 1 MB payloads, 20 vectors of 10^4-10^5 children) tiled"})
        w._keep = keep
        return w
    raise ValueError("unknown config %r" % (cfg,))


def sample(cfg, seed, threads):
    """A bounded sample of the same workload for the CPU arms: enough sessions for every thread,
    small enough that one faithful-mode pass takes a fraction of a second to a few seconds."""
    if cfg in ("cfg2", "cfg5"):
        frames = FRAMES_PER_SESSION[cfg]
        return _paired(cfg, seed, max(threads * 4, 64) if cfg == "cfg2" else max(threads * 2, 32), frames, False)
    if cfg == "cfg1":
        return make(cfg, seed, gb=0.348, nsessions=max(threads * 2, 16))
    if cfg == "cfg3":
        return make(cfg, seed, gb=0.25, nsessions=max(threads * 2, 16))
    if cfg == "cfg4":
        return make(cfg, seed, gb=0.04 * max(threads, 8))
    raise ValueError(cfg)
