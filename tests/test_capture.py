"""The zkcap capture container (include/zkcapture.h): file layout, reader checks,
and the reason it has segments — order-dependent xid pairing (lib/zk-streams.js:145 writes the
map, lib/zk-buffer.js:288-290 reads it when the reply arrives) for connections that re-use xids."""
import importlib
import os
import random
import struct

import numpy as np
import pytest

cap = importlib.import_module("node-zkstream_b200.capture")


def _phase(synth, rng, ops, first):
    """One quiescent stretch of a connection: xids 0..len(ops)-1, every request answered."""
    rq, rs = [], []
    if first:
        rq.append(synth.frame(synth.connect_request(rng)))
        rs.append(synth.frame(synth.connect_response(rng)))
    for xid, op in enumerate(ops):
        rq.append(synth.frame(synth.be32(xid) + synth.be32(op) + synth.request_body(rng, op)))
        rs.append(synth.frame(synth.resp_header(xid, 100 + xid, 0) + synth.response_body(rng, op)))
    return b"".join(rq), b"".join(rs)


def _connections(synth, seed=7):
    rng = random.Random(seed)
    a1 = _phase(synth, rng, [4, 3, 8, 12, 1], True)       # GET_DATA, EXISTS, GET_CHILDREN, GET_CHILDREN2, CREATE
    a2 = _phase(synth, rng, [3, 12, 4, 2, 5, 9], False)   # same xids, other opcodes
    a3 = _phase(synth, rng, [12, 12, 4], False)
    b1 = _phase(synth, rng, [5, 4, 3, 2], True)
    c1 = _phase(synth, rng, [], True)                     # handshake only
    return [[a1, a2, a3], b1, [c1, _phase(synth, rng, [8, 4], False)]]


def _oracle_paired(oracle, zk, c):
    n = c.nsegments
    req = oracle.decode(c.req, c.req_off, np.full(n, zk._abi.SESS_SERVER, np.uint8), c.req_hs, None)
    xt = oracle.xidtab_build(req, zk.XidTable(np.array(c.xid_base), np.array(c.xid_count)))
    rsp = oracle.decode(c.rsp, c.rsp_off, None, c.rsp_hs, xt)
    return req, rsp


def test_layout_matches_header(tmp_path, synth):
    path = str(tmp_path / "a.zkcap")
    conns = _connections(synth)
    assert cap.write_capture(path, conns) == 6
    raw = open(path, "rb").read()
    hdr = open(os.path.join(os.path.dirname(__file__), "..", "include", "zkcapture.h")).read()
    assert "128 bytes" in hdr and len(raw) % 4096 == 0
    f = struct.unpack_from("<8sII12Q", raw, 0)
    assert f[0] == b"ZKCAP\0\0\1" and f[1] == 1 and f[2] == 6 and f[-1] == len(raw)
    assert all(o % 4096 == 0 and o > 0 for o in f[5:14])
    with cap.Capture(path) as c:
        assert c.nsegments == 6
        assert list(c.conn) == [0, 0, 0, 1, 2, 2]
        assert list(c.req_hs) == [1, 0, 0, 1, 1, 0] and list(c.rsp_hs) == [1, 0, 0, 1, 1, 0]
        assert list(c.xid_count) == [6, 6, 3, 5, 1, 2]       # complete frames per request part (connect frame included)
        flat = b"".join(s[0] for conn in conns for s in (conn if isinstance(conn, list) else [conn]))
        assert bytes(c.req) == flat
        assert int(c.req_off[-1]) == len(flat) and int(c.rsp_off[-1]) == c.rsp.size


def test_reader_rejects_damaged_files(tmp_path, synth):
    path = str(tmp_path / "a.zkcap")
    cap.write_capture(path, _connections(synth))
    raw = bytearray(open(path, "rb").read())
    bad = str(tmp_path / "b.zkcap")
    open(bad, "wb").write(raw[:-4096])
    with pytest.raises(cap.CaptureError, match="truncated"):
        cap.Capture(bad)
    raw2 = bytearray(raw)
    raw2[0:5] = b"XXXXX"
    open(bad, "wb").write(raw2)
    with pytest.raises(cap.CaptureError, match="not a zkcap"):
        cap.Capture(bad)
    open(bad, "wb").write(b"short")
    with pytest.raises(cap.CaptureError):
        cap.Capture(bad)


def test_segments_make_xid_reuse_exact(tmp_path, oracle, synth, zk):
    """Connection A uses xids 0..5 three times with different opcodes.  As one flat session
    'last write wins' gives the early replies the later opcodes; cut at the quiescent points
    every reply gets the opcode that was in the map when it arrived."""
    conns = _connections(synth)
    path = str(tmp_path / "a.zkcap")
    cap.write_capture(path, conns)
    with cap.Capture(path) as c:
        req, rsp = _oracle_paired(oracle, zk, c)
        assert (rsp.status == 0).all() and (req.status == 0).all()
        ops = [(int(m) >> 8) & 0xFF for m in rsp.meta]
        first = [int(x) for x in rsp.sess_first_frame]
        assert ops[first[0] + 1:first[1]] == [4, 3, 8, 12, 1]
        assert ops[first[1]:first[2]] == [3, 12, 4, 2, 5, 9]
        assert ops[first[2]:first[3]] == [12, 12, 4]
    # the same connection as ONE segment: replies of the first phase are read with later opcodes
    flat = str(tmp_path / "flat.zkcap")
    a = conns[0]
    cap.write_capture(flat, [(b"".join(s[0] for s in a), b"".join(s[1] for s in a))])
    with cap.Capture(flat) as c:
        _, rsp = _oracle_paired(oracle, zk, c)
        ops = [(int(m) >> 8) & 0xFF for m in rsp.meta][1:6]
        assert ops != [4, 3, 8, 12, 1]


@pytest.mark.gpu
def test_gpu_replay_of_capture_file(tmp_path, ctx, oracle, synth, zk):
    shard = importlib.import_module("node-zkstream_b200.shard")
    rng = random.Random(3)
    conns = _connections(synth, seed=11)
    for k in range(40):                                   # plus ordinary single-segment connections
        conns.append(_phase(synth, rng, [rng.choice([1, 2, 3, 4, 5, 8, 9, 12]) for _ in range(rng.randrange(0, 60))], True))
    path = str(tmp_path / "g.zkcap")
    n = cap.write_capture(path, conns)
    with cap.Capture(path) as c:
        want_req, want_rsp = _oracle_paired(oracle, zk, c)
        for slice_bytes in (0, 4096):
            got = []
            c.replay(ctx, lambda first, ns, rq, rs: got.append((first, ns, rq, rs)), slice_bytes=slice_bytes)
            assert sum(g[1] for g in got) == n
            zeros = [0] * len(got)
            assert shard.merge([g[2] for g in got], zeros, zeros).diff(want_req) == []
            assert shard.merge([g[3] for g in got], zeros, zeros).diff(want_rsp) == []
        assert (want_rsp.status == 0).all()
