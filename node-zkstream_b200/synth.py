"""Seeded synthetic ZooKeeper wire traffic for tests and benchmarks.

The shapes follow BASELINE.json's configs (made concrete in SURVEY.md §8d): getData
replies, mixed request/response replay with xid pairing, tiny notification/ping/connect
frames, large znodes, plus a fuzzer that produces every malformed shape the reference's
readers can trip over.  Pure Python: meant for up to ~10^5 frames; bench.py uses the C
generator (csrc/zkgen.c) for multi-gigabyte corpora with the same frame shapes.
"""
import random
import struct

import numpy as np

from . import _abi
from .consts import OP_CODES

OP = OP_CODES


def be32(v):
    return struct.pack(">i", v)


def be64(v):
    return struct.pack(">q", v)


def jbuf(b):
    """A jute buffer/ustring as real servers write it (empty -> length 0)."""
    return be32(len(b)) + b


def frame(payload):
    return struct.pack(">I", len(payload)) + payload


def stat_bytes(rng, data_length=0, num_children=0):
    return (be64(rng.getrandbits(40)) + be64(rng.getrandbits(40)) +
            be64(1400000000000 + rng.getrandbits(36)) + be64(1400000000000 + rng.getrandbits(36)) +
            be32(rng.randrange(0, 1000)) + be32(rng.randrange(0, 1000)) + be32(rng.randrange(0, 10)) +
            be64(rng.choice([0, rng.getrandbits(56)])) + be32(data_length) + be32(num_children) +
            be64(rng.getrandbits(40)))


def rpath(rng, lo=4, hi=24):
    n = rng.randrange(lo, hi + 1)
    return b"/" + bytes(rng.choice(b"abcdefghijklmnopqrstuvwxyz0123456789-_/") for _ in range(n - 1))


def resp_header(xid, zxid, err):
    return be32(xid) + be64(zxid) + be32(err)


def connect_response(rng, passwd_len=16, read_only=True):
    p = be32(0) + be32(30000) + be64(rng.getrandbits(62)) + jbuf(bytes(rng.getrandbits(8) for _ in range(passwd_len)))
    return p + (b"\x00" if read_only else b"")


def connect_request(rng, passwd_len=16, read_only=True):
    p = (be32(0) + be64(rng.getrandbits(40)) + be32(30000) + be64(rng.getrandbits(62)) +
         jbuf(bytes(passwd_len)))
    return p + (b"\x00" if read_only else b"")


def acl_bytes(entries):
    out = be32(len(entries))
    for perms, scheme, ident in entries:
        out += be32(perms) + jbuf(scheme) + jbuf(ident)
    return out


WORLD_ACL = [(31, b"world", b"anyone")]


def request_body(rng, op, path=None, data=None):
    """Body (after xid + type) of a client request."""
    path = rpath(rng) if path is None else path
    if op in (OP["GET_CHILDREN"], OP["GET_CHILDREN2"], OP["GET_DATA"], OP["EXISTS"]):
        return jbuf(path) + bytes([rng.randrange(2)])
    if op == OP["CREATE"]:
        d = bytes(rng.getrandbits(8) for _ in range(rng.randrange(0, 64))) if data is None else data
        return jbuf(path) + jbuf(d) + acl_bytes(WORLD_ACL) + be32(rng.randrange(4))
    if op == OP["DELETE"]:
        return jbuf(path) + be32(rng.randrange(-1, 50))
    if op in (OP["GET_ACL"], OP["SYNC"]):
        return jbuf(path)
    if op == OP["SET_DATA"]:
        d = bytes(rng.getrandbits(8) for _ in range(rng.randrange(0, 64))) if data is None else data
        return jbuf(path) + jbuf(d) + be32(rng.randrange(-1, 50))
    if op == OP["SET_WATCHES"]:
        out = be64(rng.getrandbits(40))
        for _ in range(3):
            n = rng.randrange(0, 4)
            out += be32(n) + b"".join(jbuf(rpath(rng)) for _ in range(n))
        return out
    return b""   # PING, CLOSE_SESSION: header only


def response_body(rng, op, nchildren=None, data_len=None):
    """Body of an OK reply to `op`."""
    if op in (OP["GET_CHILDREN"], OP["GET_CHILDREN2"]):
        n = rng.randrange(0, 8) if nchildren is None else nchildren
        out = be32(n) + b"".join(jbuf(rpath(rng, 3, 12)[1:]) for _ in range(n))
        if op == OP["GET_CHILDREN2"]:
            out += stat_bytes(rng, 0, n)
        return out
    if op == OP["CREATE"]:
        return jbuf(rpath(rng))
    if op == OP["GET_ACL"]:
        k = rng.randrange(0, 3)
        ents = [(rng.randrange(32), b"digest", b"user:" + bytes(rng.choice(b"abcdef0123") for _ in range(8)))
                for _ in range(k)]
        return acl_bytes(ents) + stat_bytes(rng)
    if op == OP["GET_DATA"]:
        n = rng.randrange(0, 300) if data_len is None else data_len
        return jbuf(bytes(rng.getrandbits(8) for _ in range(n))) + stat_bytes(rng, n, 0)
    if op in (OP["EXISTS"], OP["SET_DATA"]):
        return stat_bytes(rng)
    return b""


def notification(rng):
    return (resp_header(-1, -1, 0) + be32(rng.randrange(1, 5)) + be32(3) + jbuf(rpath(rng, 4, 12)))


class Corpus:
    """A batch: concatenated session streams + what a caller has to say about them."""

    def __init__(self):
        self.chunks = []      # per session: bytes
        self.flags = []
        self.hs = []
        self.xidmaps = []     # per session: {xid: wire opcode}

    def add(self, data, server=False, hs=1, xidmap=None):
        self.chunks.append(bytes(data))
        self.flags.append(_abi.SESS_SERVER if server else 0)
        self.hs.append(hs)
        self.xidmaps.append(dict(xidmap or {}))
        return len(self.chunks) - 1

    @property
    def data(self):
        return b"".join(self.chunks)

    @property
    def sess_off(self):
        off = [0]
        for c in self.chunks:
            off.append(off[-1] + len(c))
        return off

    def arrays(self):
        return (np.frombuffer(self.data, dtype=np.uint8), np.array(self.sess_off, dtype=np.uint64),
                np.array(self.flags, dtype=np.uint8), np.array(self.hs, dtype=np.uint32))

    def xidtable(self):
        from ._lib import XidTable
        return XidTable.from_maps(self.xidmaps)


def cfg1_getdata(n, seed=1, payload=256):
    """BASELINE config 1: one client-side stream of `n` GET_DATA OK replies, 256-byte data + Stat
    (348-byte frames), after the ConnectResponse every real stream starts with."""
    rng = random.Random(seed)
    parts = [frame(connect_response(rng))]
    xm = {}
    for i in range(n):
        data = rng.getrandbits(8 * payload).to_bytes(payload, "little") if payload else b""
        body = resp_header(i, 0x100000000 + i, 0) + jbuf(data) + stat_bytes(rng, payload, 0)
        parts.append(frame(body))
        xm[i] = OP["GET_DATA"]
    c = Corpus()
    c.add(b"".join(parts), xidmap=xm)
    return c


MIX_OPS = ("CREATE", "GET_CHILDREN2", "SET_DATA", "EXISTS", "PING")
ERRS = (-101, -110, -103)


def session_pair(rng, nframes, ops=MIX_OPS, err_rate=0.05, max_data=1024, max_children=32):
    """One connection's two directions: (request stream, response stream, xid map)."""
    req = [frame(connect_request(rng))]
    rsp = [frame(connect_response(rng))]
    xm = {}
    xid = 0
    zxid = 0x100000000
    for _ in range(nframes):
        name = rng.choice(ops)
        op = OP[name]
        if name == "PING":
            req.append(frame(be32(-2) + be32(op)))
            rsp.append(frame(resp_header(-2, zxid, 0)))
            continue
        path = rpath(rng, 16, 64)
        data = None
        if name in ("CREATE", "SET_DATA"):
            n = rng.randrange(0, max_data + 1)
            data = rng.getrandbits(8 * n).to_bytes(n, "little") if n else b""
        req.append(frame(be32(xid) + be32(op) + request_body(rng, op, path, data)))
        zxid += 1
        if rng.random() < err_rate:
            rsp.append(frame(resp_header(xid, zxid, rng.choice(ERRS))))
        else:
            rsp.append(frame(resp_header(xid, zxid, 0) +
                             response_body(rng, op, nchildren=rng.randrange(0, max_children + 1))))
        xm[xid] = op
        xid += 1
    return b"".join(req), b"".join(rsp), xm


def cfg2_mixed(nsessions, frames_per_session, seed=2, **kw):
    """BASELINE config 2: paired request/response replay.  Sessions 0..n-1 are the client->server
    directions (ZKB_SESS_SERVER), n..2n-1 the matching server->client directions."""
    rng = random.Random(seed)
    pairs = [session_pair(rng, frames_per_session, **kw) for _ in range(nsessions)]
    c = Corpus()
    for rq, _, _ in pairs:
        c.add(rq, server=True)
    for _, rs, xm in pairs:
        c.add(rs, xidmap=xm)
    return c


def cfg3_tiny(nsessions, frames_per_session, seed=3):
    """BASELINE config 3: notifications (36-44 B), ping replies (20 B), connect frames."""
    rng = random.Random(seed)
    c = Corpus()
    for s in range(nsessions):
        if s % 10 == 9:
            c.add(frame(connect_request(rng)), server=True)
            continue
        parts = [frame(connect_response(rng))]
        for _ in range(frames_per_session):
            if rng.random() < 0.667:
                parts.append(frame(notification(rng)))
            else:
                parts.append(frame(resp_header(-2, rng.getrandbits(40), 0)))
        c.add(b"".join(parts))
    return c


def cfg4_large(nframes, seed=4, payload=1000000, children=(10000, 100000)):
    """BASELINE config 4: 1 MB getData replies and getChildren2 replies with 10^4-10^5 children."""
    rng = random.Random(seed)
    parts = [frame(connect_response(rng))]
    xm = {}
    for i in range(nframes):
        if i % 2 == 0:
            data = rng.getrandbits(8 * payload).to_bytes(payload, "little")
            parts.append(frame(resp_header(i, i + 1, 0) + jbuf(data) + stat_bytes(rng, payload, 0)))
            xm[i] = OP["GET_DATA"]
        else:
            n = rng.randrange(children[0], children[1] + 1)
            kids = b"".join(jbuf(b"n%07d" % k + bytes(rng.randrange(0, 16))) for k in range(n))
            parts.append(frame(resp_header(i, i + 1, 0) + be32(n) + kids + stat_bytes(rng, 0, n)))
            xm[i] = OP["GET_CHILDREN2"]
    c = Corpus()
    c.add(b"".join(parts), xidmap=xm)
    return c


# ---- fuzzer: every way a frame can go wrong ------------------------------------------------

ALL_OPS = sorted(OP.values())
REQ_OPS = [OP[k] for k in ("GET_CHILDREN", "GET_CHILDREN2", "CREATE", "DELETE", "GET_ACL", "SET_WATCHES",
                           "GET_DATA", "EXISTS", "SET_DATA", "SYNC", "PING", "CLOSE_SESSION", "AUTH",
                           "SET_ACL", "MULTI")]


def _mutate(rng, payload):
    """Damage a well-formed payload in one of the ways the readers react to."""
    k = rng.randrange(8)
    if k == 0 and payload:                      # truncate
        return payload[:rng.randrange(len(payload))]
    if k == 1:                                  # trailing garbage (ignored by the readers)
        return payload + bytes(rng.getrandbits(8) for _ in range(rng.randrange(1, 9)))
    if k == 2 and len(payload) >= 4:            # overwrite an aligned int with a wild value
        i = rng.randrange(0, len(payload) - 3)
        v = rng.choice([-1, -2, 0x7fffffff, -0x80000000, 0, 1, 255, 65536])
        return payload[:i] + be32(v) + payload[i + 4:]
    if k == 3 and payload:                      # flip one byte
        i = rng.randrange(len(payload))
        return payload[:i] + bytes([rng.getrandbits(8)]) + payload[i + 1:]
    if k == 4:
        return b""                              # zero-length frame: legal length, decode throws
    return payload


def fuzz_corpus(seed, nsessions=24, max_frames=40, p_mutate=0.3, p_tail=0.3, p_badlen=0.15):
    """Random sessions of both directions with damaged frames, unknown xids/opcodes/error codes,
    empty sessions, partial trailing frames and BAD_LENGTH prefixes."""
    rng = random.Random(seed)
    c = Corpus()
    for _ in range(nsessions):
        server = rng.random() < 0.4
        hs = rng.choice([1, 1, 1, 0, 2])
        nf = rng.choice([0, 0, 1, 2, rng.randrange(1, max_frames + 1)])
        parts = []
        xm = {}
        for i in range(nf):
            if i < hs:
                pl = (connect_request if server else connect_response)(
                    rng, passwd_len=rng.choice([0, 8, 16]), read_only=rng.random() < 0.5)
            elif server:
                op = rng.choice(REQ_OPS + [77])
                xid = rng.choice([i, i, -2, -8])
                pl = be32(xid) + be32(op) + request_body(rng, op)
            else:
                r = rng.random()
                if r < 0.15:
                    pl = notification(rng)
                elif r < 0.25:
                    pl = resp_header(-2, rng.getrandbits(40), 0)
                else:
                    op = rng.choice(ALL_OPS)
                    xid = i
                    if rng.random() < 0.9:
                        xm[xid] = op
                    err = rng.choice([0, 0, 0, 0, -101, -110, -4, -999, 5])
                    pl = resp_header(xid, rng.getrandbits(48), err)
                    if err == 0:
                        pl += response_body(rng, op)
            if rng.random() < p_mutate:
                pl = _mutate(rng, pl)
            parts.append(frame(pl))
        tail = b""
        r = rng.random()
        if r < p_badlen:
            tail = rng.choice([b"\x40\x00\x40\x00", b"\xff\xff\xff\xfe\x01\x02", b"\x01\x00\x00\x01",
                               b"\x80\x00\x00\x00"]) + bytes(rng.getrandbits(8) for _ in range(rng.randrange(0, 40)))
        elif r < p_badlen + p_tail:
            full = frame(notification(rng))
            tail = full[:rng.randrange(1, len(full))]
        c.add(b"".join(parts) + tail, server=server, hs=hs, xidmap=xm)
    return c


def session_corpus(seed, nsessions=12, max_frames=60):
    """Client-side (server->client) sessions for the session-level folds: replies whose zxids
    cover the signed range, notifications of every type and state (valid or not), error
    notifications, frames that fail to decode, the odd damaged tail."""
    rng = random.Random(0x5E55 + seed)
    c = Corpus()
    zx_pool = [0, 1, -1, 5, (1 << 63) - 1, -(1 << 63), 0x100000000, 0xFFFFFFFF, -0x100000000, 0x7FFFFFFF00000000]
    for _ in range(nsessions):
        parts, xm = [frame(connect_response(rng))], {}
        hs = 1
        if rng.random() < 0.15:
            parts, hs = [], 0
        nf = rng.randrange(0, max_frames)
        for i in range(nf):
            r = rng.random()
            if r < 0.35:
                t = rng.choice([1, 2, 3, 4, 4, 3, 0, 5, -1, 77])
                st = rng.choice([3, 3, 3, 3, 0, 4, 5, 6, -122, 9])
                err = rng.choice([0, 0, 0, 0, 0, -101, -4])
                body = (be32(t) + be32(st) + jbuf(rpath(rng, 1, 12))) if err == 0 else b""
                parts.append(frame(resp_header(-1, rng.choice([-1, -1, 7]), err) + body))
            elif r < 0.45:
                parts.append(frame(resp_header(-2, rng.choice(zx_pool), 0)))           # ping reply
            elif r < 0.52:
                parts.append(frame(resp_header(1000 + i, rng.choice(zx_pool), 0)))     # xid nobody sent
            elif r < 0.56:
                parts.append(frame(resp_header(i, 3, 0)[:rng.randrange(0, 15)]))       # short header
            else:
                op = rng.choice([1, 2, 3, 4, 5, 8, 9, 12])
                err = rng.choice([0, 0, 0, -101, -110, -103])
                zx = rng.choice(zx_pool) if rng.random() < 0.5 else rng.getrandbits(64) - (1 << 63)
                parts.append(frame(resp_header(i, zx, err) + (response_body(rng, op) if err == 0 else b"")))
                xm[i] = op
        tail = rng.choice([b"", b"", b"", b"\x00\x00", b"\xff\xff\xff\xff", b"\x00\x00\x00\x20abc"])
        c.add(b"".join(parts) + tail, hs=hs, xidmap=xm)
    return c


# ---- benchmark-scale corpora from the C generator (csrc/zkgen.c) --------------------------

_gen = None


def _genlib():
    global _gen
    if _gen is None:
        import ctypes as C
        import os
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libzkgen.so")
        if not os.path.exists(path):
            raise ImportError("%s is missing: run `python node-zkstream_b200/build.py`" % path)
        g = C.CDLL(path)
        P = C.c_void_p
        g.zkgen_cfg2_sizes.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, P, P, C.c_int]
        g.zkgen_cfg2_fill.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, P, P, P, P, P, C.c_int]
        g.zkgen_cfg1.restype = C.c_uint64
        g.zkgen_cfg1.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, P]
        g.zkgen_cfg3.restype = C.c_uint64
        g.zkgen_cfg3.argtypes = [C.c_uint64, C.c_uint64, P]
        _gen = g
    return _gen


class PairedCorpus:
    """Config-2 replay: `req`/`rsp` byte arrays with per-session offsets, and the dense xid table
    (`ops`, [nsessions * frames] wire opcodes, OP_NONE past each session's last xid)."""

    def __init__(self, nsessions, frames):
        self.nsessions, self.frames = nsessions, frames

    def xidtable(self):
        from ._lib import XidTable
        n = self.nsessions
        return XidTable(np.zeros(n, np.int32), np.full(n, self.frames, np.uint32), self.ops)


def native_cfg2(seed, nsessions, frames, nthreads=None, alloc=None):
    """Paired request/response replay (config 2) from the C generator.  `alloc(nbytes)` may
    return a writable uint8 array to fill (e.g. pinned memory); default is numpy."""
    import os
    g = _genlib()
    nthreads = nthreads or os.cpu_count() or 1
    alloc = alloc or (lambda n: np.empty(n, dtype=np.uint8))
    pc = PairedCorpus(nsessions, frames)
    pc.req_off = np.zeros(nsessions + 1, dtype=np.uint64)
    pc.rsp_off = np.zeros(nsessions + 1, dtype=np.uint64)
    g.zkgen_cfg2_sizes(seed, nsessions, frames, pc.req_off.ctypes.data, pc.rsp_off.ctypes.data, nthreads)
    pc.req = alloc(int(pc.req_off[-1]))
    pc.rsp = alloc(int(pc.rsp_off[-1]))
    pc.ops = np.full(nsessions * frames, _abi.OP_NONE, dtype=np.int8)
    g.zkgen_cfg2_fill(seed, nsessions, frames, pc.req_off.ctypes.data, pc.rsp_off.ctypes.data,
                      pc.req.ctypes.data, pc.rsp.ctypes.data, pc.ops.ctypes.data, nthreads)
    pc.req_flags = np.full(nsessions, _abi.SESS_SERVER, dtype=np.uint8)
    pc.rsp_flags = np.zeros(nsessions, dtype=np.uint8)
    return pc


def native_cfg1(seed, n, payload=256):
    g = _genlib()
    size = g.zkgen_cfg1(seed, n, payload, None)
    out = np.empty(size, dtype=np.uint8)
    g.zkgen_cfg1(seed, n, payload, out.ctypes.data)
    return out


def native_cfg3(seed, n):
    g = _genlib()
    size = g.zkgen_cfg3(seed, n, None)
    out = np.empty(size, dtype=np.uint8)
    g.zkgen_cfg3(seed, n, out.ctypes.data)
    return out
