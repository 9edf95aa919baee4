"""GPU: the CUDA codec against the committed reference fixtures, the host-side mirror of the
reference interface (ZKDecodeStream / ZKEncodeStream / ZKBuffer) and the device-pointer ABI."""
import base64
import importlib
import json
import os
import random

import numpy as np
import pytest

from _canon import materialised_events, zk
from test_oracle_golden import CAPTURE1, DECODED1, unjson

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _load(name):
    with open(os.path.join(GOLD, name)) as f:
        return json.load(f)


# ---- committed fixtures (what the reference's JavaScript produced) ----------------------

def test_cuda_decode_matches_reference_fixtures(ctx, synth):
    g = _load("ref_decode.json")
    c = synth.Corpus()
    for s in g["sessions"]:
        c.add(bytes.fromhex(s["hex"]), server=s["isServer"], hs=s["hs"],
              xidmap={int(k): v for k, v in s["xidMap"].items()})
    d, off, fl, hs = c.arrays()
    cols = ctx.decode_host(d, off, fl, hs, c.xidtable())
    for i, s in enumerate(g["sessions"]):
        assert materialised_events(cols, c.data, i) == s["events"], (s["corpus"], i)
        assert int(off[i + 1]) - int(cols.sess_consumed[i]) == s["used"], (s["corpus"], i)


def test_cuda_encode_matches_reference_fixtures(ctx):
    g = _load("ref_encode.json")
    for p in g["packets"]:
        pkt = unjson(p["pkt"])
        data, off, st = ctx.encode_host(zk.RequestBatch.from_packets([pkt], p["handshaking"], p["isServer"]))
        if "frame" in p["result"]:
            assert st[0] == 0 and bytes(data).hex() == p["result"]["frame"], pkt
        else:
            assert st[0] == 1 and off[1] == 0


# ---- encode parity at size -----------------------------------------------------------------

def random_packets(rng, n):
    pkts = []
    for i in range(n):
        op = rng.choice(["GET_DATA", "EXISTS", "GET_CHILDREN", "GET_CHILDREN2", "CREATE", "DELETE", "SET_DATA",
                         "SYNC", "GET_ACL", "PING", "CLOSE_SESSION", "SET_WATCHES", "SET_ACL"])
        path = "/" + "".join(rng.choice("abcxyz/09é") for _ in range(rng.randrange(0, 40)))
        p = dict(xid=i - 3, opcode=op)
        if op in ("GET_DATA", "EXISTS", "GET_CHILDREN", "GET_CHILDREN2"):
            p.update(path=path, watch=rng.random() < 0.5)
        elif op == "CREATE":
            p.update(path=path, data=bytes(rng.getrandbits(8) for _ in range(rng.choice([0, 1, 3, 64, 1500]))),
                     acl=[dict(id=dict(scheme="world", id="anyone"), perms=["read", "Write"])] * rng.randrange(0, 3),
                     flags=rng.choice([[], ["EPHEMERAL"], ["SEQUENTIAL", "EPHEMERAL"]]))
        elif op == "DELETE":
            p.update(path=path, version=rng.randrange(-1, 9))
        elif op == "SET_DATA":
            p.update(path=path, data=bytes(rng.getrandbits(8) for _ in range(rng.randrange(0, 700))),
                     version=rng.randrange(-1, 9))
        elif op in ("SYNC", "GET_ACL", "SET_ACL"):
            p.update(path=path)
        elif op == "SET_WATCHES":
            p.update(relZxid=bytes(rng.getrandbits(8) for _ in range(8)),
                     events=dict(dataChanged=[path] * rng.randrange(0, 3), createdOrDestroyed=[],
                                 childrenChanged=["/c", ""][: rng.randrange(0, 3)]))
        pkts.append(p)
    return pkts


@pytest.mark.parametrize("seed,n", [(1, 1), (2, 37), (3, 5000)])
def test_encode_parity(ctx, oracle, seed, n):
    batch = zk.RequestBatch.from_packets(random_packets(random.Random(seed), n))
    gd, go, gs = ctx.encode_host(batch)
    od, oo, os_ = oracle.encode(batch)
    assert np.array_equal(go, oo) and np.array_equal(gs, os_)
    assert np.array_equal(gd, od)


def test_encode_empty_batch(ctx, oracle):
    batch = zk.RequestBatch.from_packets([])
    gd, go, gs = ctx.encode_host(batch)
    assert gd.size == 0 and list(go) == [0] and gs.size == 0


def test_encode_decode_roundtrip_on_gpu(ctx):
    """encode -> decode -> same packets (a property that holds at any size)."""
    pkts = [p for p in random_packets(random.Random(8), 3000) if p["opcode"] not in ("PING", "CLOSE_SESSION", "SET_ACL")]
    data, off, st = ctx.encode_host(zk.RequestBatch.from_packets(pkts))
    assert (st == 0).all()
    cols = ctx.decode_host(data, [0, data.size], [1], [0], None)
    assert cols.n_frames == len(pkts) and (cols.status == 0).all()
    got = [p for k, p in zk.Materialiser(cols, bytes(data)).events() if k == "pkt"]
    for a, b in zip(got, pkts):
        assert a["xid"] == b["xid"] and a["opcode"] == b["opcode"]
        for k in ("path", "watch", "version", "data"):
            if k in b:
                assert a[k] == b[k], (k, b)
        if "flags" in b:      # flags come back in CREATE_FLAGS order (zk-buffer.js:154-158)
            assert sorted(a["flags"]) == sorted(b["flags"])
        if b["opcode"] == "SET_WATCHES":
            assert a["relZxid"] == b["relZxid"]
            assert a["events"]["dataChanged"] == b["events"]["dataChanged"]
            assert a["events"]["childrenChanged"] == b["events"]["childrenChanged"]


# ---- the reference's own offline test, ported ---------------------------------------------

def test_streams_test_js_decode_capture_1(ctx):
    """test/streams.test.js:58-82 with ZKBuffer from this package."""
    xidMap = {}
    for i, (direction, b64) in enumerate(CAPTURE1):
        buf = base64.b64decode(b64)
        zkb = zk.ZKBuffer({"buffer": buf, "ctx": ctx})
        length = zkb.readInt()
        assert length == len(zkb.remainder())
        if i == 0:
            pkt = zkb.readConnectRequest()
        elif i == 1:
            pkt = zkb.readConnectResponse()
        elif direction == "send":
            pkt = zkb.readRequest()
            xidMap[pkt["xid"]] = pkt["opcode"]
        else:
            pkt = zkb.readResponse(xidMap)
        assert pkt == DECODED1[i]


def test_zkbuffer_reencodes_capture(ctx):
    """writeLengthPrefixed(writeRequest) reproduces the captured request byte for byte."""
    want = base64.b64decode(CAPTURE1[2][1])
    zkb = zk.ZKBuffer({"ctx": ctx})
    zkb.writeLengthPrefixed(lambda sub: sub.writeRequest(dict(xid=1, opcode="GET_CHILDREN", path="/", watch=False)))
    assert zkb.toBuffer() == want
    zkb = zk.ZKBuffer({"ctx": ctx})
    with pytest.raises(Exception, match="Unsupported opcode SET_ACL"):
        zkb.writeRequest(dict(xid=1, opcode="SET_ACL", path="/"))


def test_zkbuffer_reader_errors(ctx):
    zkb = zk.ZKBuffer({"buffer": bytes.fromhex("00000001000000000000051700000000"), "ctx": ctx})
    with pytest.raises(AssertionError, match="reply packet must match a request"):
        zkb.readResponse({})
    zkb = zk.ZKBuffer({"buffer": bytes.fromhex("0000000100000000000005"), "ctx": ctx})
    with pytest.raises(AssertionError):
        zkb.readResponse({1: "PING"})
    zkb = zk.ZKBuffer({"buffer": bytes.fromhex("000000010000000b"), "ctx": ctx})
    with pytest.raises(Exception, match="Unsupported opcode PING"):
        zkb.readRequest()


class FakeFsm:
    """The duck-typed fsm of test/nasty.test.js:298-310, with a switchable state."""

    def __init__(self, state="handshaking"):
        self.state = state
        self.zcf_xidMap = {}

    def isInState(self, st):
        return st == self.state


def test_decode_stream_like_the_reference(ctx):
    b = [base64.b64decode(x) for _, x in CAPTURE1]
    fsm = FakeFsm()
    dec = zk.ZKDecodeStream({"fsm": fsm, "ctx": ctx})
    got, errs = [], []
    dec.on("error", errs.append)
    dec.on("readable", lambda: got.append(dec.read()))
    dec.write(b[1][:9])                 # a partial frame stays buffered
    assert got == [] and dec.zd_used == 9
    dec.write(b[1][9:])
    assert got == [DECODED1[1]] and dec.zd_used == 0
    fsm.state = "connected"
    fsm.zcf_xidMap[1] = "GET_CHILDREN"
    dec.write(b[3] + b"\x00\x00")       # a reply plus the start of the next frame
    assert got[1] == DECODED1[3] and dec.zd_used == 2
    dec.write(bytes.fromhex("0010" "00000002" "0000000000000009" "00000000"))   # xid 2 was never sent
    assert errs and errs[0].code == "BAD_DECODE"
    assert errs[0].message == ("BAD_DECODE: Failed to decode Response: AssertionError: "
                               "reply packet must match a request")
    dec.write(bytes.fromhex("fffffffe0102"))
    assert errs[-1].code == "BAD_LENGTH" and errs[-1].message == "BAD_LENGTH: Invalid ZK packet length"
    n = len(got)
    dec.write(b[3])                     # nothing more is decoded after BAD_LENGTH
    assert len(got) == n


def test_decode_stream_server_side_batch_mode(ctx, synth):
    """Replay use: one big write, thousands of frames per kernel launch."""
    c = synth.cfg2_mixed(1, 2000, seed=4)
    fsm = FakeFsm("connected")
    dec = zk.ZKDecodeStream({"fsm": fsm, "isServer": True, "ctx": ctx, "batchBytes": 1 << 30, "handshakeFrames": 1})
    got, errs = [], []
    dec.on("data", got.append)
    dec.on("error", errs.append)
    dec.write(c.chunks[0])
    assert got == []
    dec.end()
    assert len(got) + len(errs) == 2001
    assert got[0]["protocolVersion"] == 0 and "passwd" in got[0]
    assert all(e.message.endswith("Unsupported opcode PING") for e in errs)     # readRequest cannot read PING


def test_encode_stream_like_the_reference(ctx):
    fsm = FakeFsm()
    enc = zk.ZKEncodeStream({"fsm": fsm, "ctx": ctx})
    out = []
    enc.on("data", out.append)
    enc.write(dict(protocolVersion=0, lastZxidSeen=bytes(8), timeOut=30000, sessionId=bytes(8), passwd=bytes(16)))
    cap = base64.b64decode(CAPTURE1[0][1])                      # zkCli's capture ends in a readOnly byte
    assert out[0][4:] == cap[4:-1] and out[0][:4] == bytes([0, 0, 0, len(cap) - 5])
    assert fsm.zcf_xidMap == {}
    fsm.state = "connected"
    enc.write(dict(xid=1, opcode="GET_CHILDREN", path="/", watch=False))
    assert out[1] == base64.b64decode(CAPTURE1[2][1])
    assert fsm.zcf_xidMap == {1: "GET_CHILDREN"}                # zk-streams.js:145
    with pytest.raises(AssertionError):
        enc.write(dict(opcode="PING"))                          # mod_assert.number(pkt.xid, 'xid')
    srv = zk.ZKEncodeStream({"fsm": FakeFsm(), "isServer": True, "ctx": ctx})
    res = []
    srv.on("data", res.append)
    srv.write(dict(protocolVersion=1, timeOut=30000, sessionId=bytes(8), passwd=bytes(8)))   # nasty.test.js:325-331
    assert res[0].hex() == "0000001c" + "00000001" + "00007530" + "00" * 8 + "00000008" + "00" * 8


# ---- device-pointer entry points -------------------------------------------------------------

def test_decode_dev_matches_host_variant(ctx, oracle, synth):
    import torch
    dev = importlib.import_module("node-zkstream_b200.device")
    c = synth.fuzz_corpus(77, nsessions=300, max_frames=120)
    d, off, fl, hs = c.arrays()
    xt = c.xidtable()
    want = oracle.decode(d, off, fl, hs, xt)
    d_bytes = dev.to_device(d, "cuda:0", pad=64)
    ds = dev.DeviceSessions(d.size, off, fl, hs, "cuda:0")
    dx = dev.DeviceXidTable(xt, "cuda:0")
    # too small on purpose: the call must say how many rows it needs and write nothing out of bounds
    dc = dev.DeviceColumns(ds.n, 8, 1, 1, 1, "cuda:0")
    rc = dev.decode_dev(ctx, d_bytes, d.size, ds, dx, dc)
    assert rc == zk._abi.EOVERFLOW and dc.struct.n_frames == want.n_frames
    dc = dev.DeviceColumns(ds.n, want.n_frames, 1, 1, 1, "cuda:0")
    rc = dev.decode_dev(ctx, d_bytes, d.size, ds, dx, dc)
    assert rc == zk._abi.EOVERFLOW
    assert (dc.struct.n_stat, dc.struct.n_buf, dc.struct.n_acl) == (want.n_stat, want.n_buf, want.n_acl)
    dc = dev.decode_dev_autosize(ctx, d_bytes, d.size, ds, dx, device="cuda:0")
    assert dc.to_host().diff(want) == []


def test_bad_arguments(ctx):
    with pytest.raises(zk.ZKBatchError) as e:
        ctx.decode_host(b"\x00" * 10, [0, 4, 3, 10])      # offsets not monotone
    assert e.value.code == zk._abi.EINVAL
    with pytest.raises(zk.ZKBatchError):
        ctx.decode_host(b"\x00" * 10, [0, 9])             # does not cover the batch


def test_large_cfg1_properties(ctx, synth):
    """Full-size config 1 (1M getData replies, 348 MB): properties that need no oracle."""
    data = synth.native_cfg1(5, 1000000, 256)
    xt = zk.XidTable([0], [1000000], np.full(1000000, 4, np.int8))
    cols = ctx.decode_host(data, None, None, None, xt, copy=False)
    assert cols.n_frames == 1000001 and cols.n_stat == 1000000 and cols.n_buf == 1000001
    assert (cols.status == 0).all()
    assert np.array_equal(cols.xid[1:], np.arange(1000000, dtype=np.int32))
    assert np.array_equal(cols.zxid[1:], 0x100000000 + np.arange(1000000, dtype=np.int64))
    assert (cols.frame_len[1:] == 344).all() and (cols.buf_len[1:] == 256).all()
    assert np.array_equal(cols.buf_off[1:], cols.frame_off[1:] + 24)
    assert (cols.dataLength == 256).all() and np.array_equal(cols.stat_frame, np.arange(1, 1000001, dtype=np.uint32))
    assert int(cols.sess_consumed[0]) == data.size


def test_replay_host_pipelined_matches_oracle(ctx, oracle, synth):
    """zkb_replay_host: slices flow through upload/decode/download streams; put back together the
    slices must equal one oracle decode of each direction (xid table derived from the requests)."""
    shard = importlib.import_module("node-zkstream_b200.shard")
    pc = synth.native_cfg2(99, 37, 900)
    n = pc.nsessions
    layout = zk.XidTable(np.zeros(n, np.int32), np.full(n, pc.frames, np.uint32))
    want_req = oracle.decode(pc.req, pc.req_off, pc.req_flags, None, None)
    want_rsp = oracle.decode(pc.rsp, pc.rsp_off, pc.rsp_flags, None, oracle.xidtab_build(want_req, layout))
    for slice_bytes in (0, 1 << 20, 300000):          # one slice, a few, many (forces buffer reuse)
        got = []
        ctx.replay_host(pc.req, pc.req_off, pc.rsp, pc.rsp_off, layout.base, layout.count,
                        lambda first, ns, rq, rs: got.append((first, ns, rq, rs)), slice_bytes=slice_bytes)
        assert [g[0] for g in got] == sorted(g[0] for g in got) and sum(g[1] for g in got) == n
        if slice_bytes:
            assert len(got) > 1
        zeros = [0] * len(got)
        req = shard.merge([g[2] for g in got], zeros, zeros)
        rsp = shard.merge([g[3] for g in got], zeros, zeros)
        assert req.diff(want_req) == [], slice_bytes
        assert rsp.diff(want_rsp) == [], slice_bytes
    assert (want_rsp.status == 0).all()


def test_replay_host_compact_results_restore_to_the_full_columns(zk, oracle, synth):
    """zkb_ctx_set_replay_compact: the replay leaves frame_off, session, the *_frame columns and buf_off
    out of the download (buf_rel instead); Columns.restore_derived must rebuild exactly what the full
    replay delivers, slice by slice, on clean and on damaged traffic."""
    pc = synth.native_cfg2(77, 21, 1200)
    n = pc.nsessions
    rsp = pc.rsp.copy()
    rsp[int(pc.rsp_off[3]) + 200] ^= 0x40             # some damage: whatever it breaks, both modes must agree
    layout = zk.XidTable(np.zeros(n, np.int32), np.full(n, pc.frames, np.uint32))
    c = zk.Context(0)
    try:
        for slice_bytes in (0, 400000):
            full, comp = [], []
            c.set_replay_compact(False)
            c.replay_host(pc.req, pc.req_off, rsp, pc.rsp_off, layout.base, layout.count,
                          lambda first, ns, rq, rs: full.append((first, ns, rq, rs)), slice_bytes=slice_bytes)
            c.set_replay_compact(True)
            c.replay_host(pc.req, pc.req_off, rsp, pc.rsp_off, layout.base, layout.count,
                          lambda first, ns, rq, rs: comp.append((first, ns, rq, rs)), slice_bytes=slice_bytes)
            assert len(full) == len(comp) and (slice_bytes == 0 or len(full) > 1)
            for (f0, n0, frq, frs), (f1, n1, crq, crs) in zip(full, comp):
                assert (f0, n0) == (f1, n1)
                assert crq.frame_off.size == 0 and crq.buf_off.size == 0 and crq.buf_rel.size == crq.n_buf
                assert crq.nbytes_columns() < 0.75 * frq.nbytes_columns()
                assert crq.restore_derived(pc.req_off[f0:f0 + n0 + 1], f0).diff(frq) == []
                assert crs.restore_derived(pc.rsp_off[f0:f0 + n0 + 1], f0).diff(frs) == []
    finally:
        c.close()


def test_reencode_round_trip_at_scale(ctx, synth):
    """Config 5's encode half: ~3.2 M requests (830 MB) decoded on the GPU, turned back into
    encode descriptors (RequestBatch.from_columns) and encoded on the GPU must reproduce the wire
    bytes — up to the two places where the reference's writers differ from what a server-side
    peer may send: empty buffers are written as length -1 (jute-buffer.js:127-130) and the
    ConnectRequest's trailing readOnly byte is not written (zk-buffer.js:32-39)."""
    pc = synth.native_cfg2(31, 160, 20000)
    cols = ctx.decode_host(pc.req, pc.req_off, pc.req_flags, None, None)
    rb, idx = zk.RequestBatch.from_columns(cols, pc.req)
    assert rb.n == cols.n_frames == 160 * 20001
    enc, fo, st = ctx.encode_host(rb)
    assert (st == 0).all()
    want = pc.req.copy()
    # empty buffers: the generator writes length 0, the reference's writer -1
    empty = np.nonzero(cols.buf_len == 0)[0]
    pos = cols.buf_off[empty].astype(np.int64) - 4
    was_zero = (want[pos] == 0) & (want[pos + 1] == 0) & (want[pos + 2] == 0) & (want[pos + 3] == 0)
    for k in range(4):
        want[pos[was_zero] + k] = 0xFF
    # connect frames: drop the readOnly byte, shorten the prefix by one
    first = cols.sess_first_frame[:-1].astype(np.int64)
    assert (cols.kind[first] == zk._abi.KIND_CONNECT_REQUEST).all()
    foff = cols.frame_off[first].astype(np.int64)
    flen = cols.frame_len[first].astype(np.int64)
    for k in range(4):
        want[foff + k] = ((flen - 1) >> (8 * (3 - k))) & 0xFF
    want = np.delete(want, foff + 4 + flen - 1)
    assert enc.size == want.size
    assert np.array_equal(enc, want)
    assert int(fo[-1]) == enc.size


def test_encode_dev_matches_host_variant(ctx, oracle):
    """zkb_encode_dev (descriptors and output in HBM) against zkb_encode_host and the oracle,
    including the too-small-output answer."""
    dev = importlib.import_module("node-zkstream_b200.device")
    rng = random.Random(12)
    pkts = []
    for i in range(3000):
        k = rng.randrange(6)
        if k == 0:
            pkts.append(dict(xid=i, opcode="SET_DATA", path="/p/%d" % i, data=bytes(rng.randrange(0, 300)), version=i))
        elif k == 1:
            pkts.append(dict(xid=i, opcode="CREATE", path="/c/%d" % i, data=bytes(rng.randrange(0, 40)),
                             acl=[dict(id=dict(scheme="world", id="anyone"), perms=["read", "write"])] * rng.randrange(0, 3),
                             flags=["EPHEMERAL"] if i % 2 else []))
        elif k == 2:
            pkts.append(dict(xid=i, opcode="GET_CHILDREN2", path="/", watch=bool(i & 1)))
        elif k == 3:
            pkts.append(dict(xid=-8, opcode="SET_WATCHES", relZxid=bytes(8),
                             events=dict(dataChanged=["/a%d" % j for j in range(rng.randrange(0, 4))], childrenChanged=["/b"])))
        elif k == 4:
            pkts.append(dict(xid=-2, opcode="PING"))
        else:
            pkts.append(dict(xid=i, opcode="SET_ACL", path="/x"))          # unsupported: status 1, no bytes
    batch = zk.RequestBatch.from_packets(pkts)
    hd, ho, hs = ctx.encode_host(batch)
    od, oo, os_ = oracle.encode(batch)
    assert np.array_equal(hd, od) and np.array_equal(ho, oo) and np.array_equal(hs, os_)
    dr = dev.DeviceRequests(batch)
    out, fo, st, n, rc = dev.encode_dev(ctx, dr, hd.size)
    assert rc == 0 and n == hd.size
    assert np.array_equal(out[:n].cpu().numpy(), hd)
    assert np.array_equal(fo.cpu().numpy().view(np.uint64), ho) and np.array_equal(st[:len(pkts)].cpu().numpy(), hs)
    _, _, _, need, rc = dev.encode_dev(ctx, dr, hd.size - 1)
    assert rc == zk._abi.EOVERFLOW and need == hd.size


def test_replay_host_with_damaged_frames(ctx, oracle, synth):
    """The pipelined replay with a few lying child counts and a truncated session in the reply
    streams: slices that contain them take the two-pass route, the result is still the oracle's."""
    shard = importlib.import_module("node-zkstream_b200.shard")
    pc = synth.native_cfg2(77, 24, 1500)
    n = pc.nsessions
    layout = zk.XidTable(np.zeros(n, np.int32), np.full(n, pc.frames, np.uint32))
    want_req = oracle.decode(pc.req, pc.req_off, pc.req_flags, None, None)
    xt = oracle.xidtab_build(want_req, layout)
    clean = oracle.decode(pc.rsp, pc.rsp_off, pc.rsp_flags, None, xt)
    rsp = pc.rsp.copy()
    gc2 = np.nonzero((clean.opcode == 12) & (clean.err == 0) & (clean.a0 > 1))[0]
    for f in gc2[:: max(1, len(gc2) // 7)][:7]:
        pos = int(clean.frame_off[f]) + 4 + 16
        rsp[pos:pos + 4] = np.frombuffer((int(clean.a0[f]) + 3).to_bytes(4, "big"), dtype=np.uint8)
    rsp[int(pc.rsp_off[5]) + 100000:int(pc.rsp_off[6])] = 0xEE       # session 5: garbage tail -> BAD_LENGTH
    want_rsp = oracle.decode(rsp, pc.rsp_off, pc.rsp_flags, None, xt)
    assert int(want_rsp.sess_bad_length_at[5]) != zk._abi.NO_OFFSET
    before = ctx.row_mispredicts
    for slice_bytes in (0, 1 << 20):
        got = []
        ctx.replay_host(pc.req, pc.req_off, rsp, pc.rsp_off, layout.base, layout.count,
                        lambda first, ns, rq, rs: got.append((first, ns, rq, rs)), slice_bytes=slice_bytes)
        zeros = [0] * len(got)
        assert shard.merge([g[2] for g in got], zeros, zeros).diff(want_req) == []
        assert shard.merge([g[3] for g in got], zeros, zeros).diff(want_rsp) == []
    assert ctx.row_mispredicts > before
