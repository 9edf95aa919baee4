"""GPU parity: libzkbatch.so (through its C ABI, host-buffer entry point) against the CPU
oracle on the same seeded bytes — every column bit-exact."""
import base64

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CAPTURE1 = [  # reference test/streams.test.js:21-27
    ("send", "AAAALQAAAAAAAAAAAAAAAAAAdTAAAAAAAAAAAAAAABAAAAAAAAAAAAAAAAAAAAAAAA=="),
    ("recv", "AAAAJQAAAAAAAHUwAVWjqFbbAAAAAAAQh19uvwgo25o9B6hUkSvqKQA="),
    ("send", "AAAADgAAAAEAAAAIAAAAAS8A"),
    ("recv", "AAAAKAAAAAEAAAAAAAAFFwAAAAAAAAACAAAACXpvb2tlZXBlcgAAAANmb28="),
]


def both(ctx, oracle, corpus, xt="auto"):
    d, off, fl, hs = corpus.arrays()
    if xt == "auto":
        xt = corpus.xidtable()
    g = ctx.decode_host(d, off, fl, hs, xt)
    o = oracle.decode(d, off, fl, hs, xt)
    return g, o


def test_capture1(ctx, oracle, synth):
    b = [base64.b64decode(x) for _, x in CAPTURE1]
    c = synth.Corpus()
    c.add(b[0] + b[2], server=True)
    c.add(b[1] + b[3], xidmap={1: 8})
    g, o = both(ctx, oracle, c)
    assert g.diff(o) == []
    assert g.n_frames == 4 and g.n_buf == 5
    assert g.a0[3] == 2 and g.zxid[3] == 0x517
    assert bytes(c.data[int(g.buf_off[3]):int(g.buf_off[3]) + int(g.buf_len[3])]) == b"zookeeper"


@pytest.mark.parametrize("seed", range(40))
def test_fuzz(ctx, oracle, synth, seed):
    g, o = both(ctx, oracle, synth.fuzz_corpus(seed))
    assert g.diff(o) == []


@pytest.mark.parametrize("seed", range(4))
def test_fuzz_big(ctx, oracle, synth, seed):
    g, o = both(ctx, oracle, synth.fuzz_corpus(1000 + seed, nsessions=400, max_frames=300))
    assert g.diff(o) == []
    assert g.n_frames > 1000


def test_cfg1_getdata(ctx, oracle, synth):
    g, o = both(ctx, oracle, synth.cfg1_getdata(20000))
    assert g.diff(o) == []
    assert g.n_frames == 20001 and g.n_stat == 20000 and g.n_buf == 20001
    assert (g.status == 0).all()


def test_cfg2_mixed(ctx, oracle, synth):
    g, o = both(ctx, oracle, synth.cfg2_mixed(16, 1500))
    assert g.diff(o) == []


def test_cfg2_xidtab_from_requests(ctx, oracle, synth, zk):
    """Build the response side's xid table from the decoded request side (zk-streams.js:145)."""
    c = synth.cfg2_mixed(8, 400, seed=9)
    n = 8
    d, off, fl, hs = c.arrays()
    req = ctx.decode_host(d[: int(off[n])], off[: n + 1], fl[:n], hs[:n], None)
    layout = zk.XidTable(np.zeros(n, np.int32), np.full(n, 400, np.uint32))
    xt_g = ctx.xidtab_build_host(req, layout)
    xt_o = oracle.xidtab_build(oracle.decode(d[: int(off[n])], off[: n + 1], fl[:n], hs[:n], None), layout)
    assert np.array_equal(xt_g.ops, xt_o.ops)
    rsp = d[int(off[n]):]
    roff = off[n:] - off[n]
    g = ctx.decode_host(rsp, roff, fl[n:], hs[n:], xt_g)
    o = oracle.decode(rsp, roff, fl[n:], hs[n:], xt_o)
    assert g.diff(o) == []
    # every reply that names a request decodes (PING replies use the special xid)
    assert (g.status == 0).all()


@pytest.mark.parametrize("seed", range(4))
def test_xidtab_reused_xids_last_request_wins(ctx, oracle, synth, zk, seed):
    """zcf_xidMap[xid] = opcode overwrites (zk-streams.js:145): when requests of a connection share an
    xid the last one in stream order names the reply's opcode.  The build stores without atomics first
    and falls back to the ordered build when a slot was written twice; both routes must agree with the
    oracle - reused xids with different and with equal opcodes, xids outside the layout, sparse tables."""
    import random
    rng = random.Random(seed)
    ops = [synth.OP[k] for k in ("CREATE", "GET_DATA", "EXISTS", "SET_DATA", "GET_CHILDREN2", "DELETE", "SYNC")]
    c = synth.Corpus()
    nsess = 6
    for s in range(nsess):
        parts = [synth.frame(synth.connect_request(rng))]
        reuse = seed > 0 and s % 2 == 0                      # seed 0: no xid is ever reused (first route alone)
        for i in range(rng.randrange(50, 400)):
            xid = rng.randrange(0, 120) if reuse else 3 * i
            op = rng.choice(ops) if seed != 2 else ops[1]     # seed 2: reused xids always carry the same opcode
            parts.append(synth.frame(synth.be32(xid) + synth.be32(op) + synth.request_body(rng, op)))
        c.add(b"".join(parts), server=True)
    d, off, fl, hs = c.arrays()
    req = ctx.decode_host(d, off, fl, hs, None)
    req_o = oracle.decode(d, off, fl, hs, None)
    assert req.diff(req_o) == []
    layout = zk.XidTable(np.array([0, 5, 0, -3, 10, 0], np.int32), np.array([1300, 100, 64, 200, 0, 1200], np.uint32))
    xt_g = ctx.xidtab_build_host(req, layout)
    xt_o = oracle.xidtab_build(req_o, layout)
    assert np.array_equal(xt_g.ops, xt_o.ops)
    assert (xt_g.ops != -128).any()


def test_cfg3_tiny(ctx, oracle, synth):
    g, o = both(ctx, oracle, synth.cfg3_tiny(50, 800))
    assert g.diff(o) == []


def test_cfg4_large(ctx, oracle, synth):
    g, o = both(ctx, oracle, synth.cfg4_large(6, payload=300000, children=(3000, 9000)))
    assert g.diff(o) == []


def test_empty_and_ragged(ctx, oracle, synth):
    c = synth.Corpus()
    g, o = both(ctx, oracle, c, xt=None)          # zero sessions, zero bytes
    assert g.diff(o) == [] and g.n_frames == 0
    c = synth.Corpus()
    c.add(b""); c.add(b""); c.add(b"\x00"); c.add(b""); c.add(b"\x00\x00\x00"); c.add(b"")
    g, o = both(ctx, oracle, c, xt=None)
    assert g.diff(o) == [] and g.n_frames == 0
    assert list(g.sess_consumed) == [0, 0, 0, 1, 1, 4]


def test_bad_lengths(ctx, oracle, synth):
    """reference test/nasty.test.js:125-126,148,170"""
    c = synth.Corpus()
    c.add(bytes.fromhex("40004000"))
    c.add(bytes.fromhex("000000000102"))
    c.add(bytes.fromhex("fffffffe0102"))
    c.add(bytes.fromhex("01000000") + bytes(16))          # exactly 16 MiB: legal, incomplete
    c.add(bytes.fromhex("01000001") + bytes(16))          # 16 MiB + 1: BAD_LENGTH
    g, o = both(ctx, oracle, c, xt=None)
    assert g.diff(o) == []
    NO = 0xFFFFFFFFFFFFFFFF
    assert list(g.sess_bad_length_at) == [0, NO, 10, NO, 36]
    assert g.n_frames == 1 and g.status[0] == 1          # zero-length frame: BAD_DECODE


def test_adversarial_split(ctx, oracle, synth):
    """Payloads that look like frame headers at every offset must not fool the splitter."""
    import random
    rng = random.Random(5)
    fake = synth.frame(synth.resp_header(7, 9, 0) + synth.stat_bytes(rng))
    parts = [synth.frame(synth.connect_response(rng))]
    xm = {}
    for i in range(3000):
        data = (fake * rng.randrange(1, 40))[rng.randrange(0, 5):]
        parts.append(synth.frame(synth.resp_header(i, i, 0) + synth.jbuf(data) + synth.stat_bytes(rng)))
        xm[i] = 4
    c = synth.Corpus()
    c.add(b"".join(parts), xidmap=xm)
    before = ctx.split_fixups
    g, o = both(ctx, oracle, c)
    assert g.diff(o) == []
    assert g.n_frames == 3001
    print("split fixups used:", ctx.split_fixups - before)


def test_native_cfg2_parity_at_size(ctx, oracle, synth, zk):
    """~190 MB of the bench workload (C generator), both directions, every column vs the oracle."""
    pc = synth.native_cfg2(5, 24, 20000)
    n = pc.nsessions
    req_g = ctx.decode_host(pc.req, pc.req_off, pc.req_flags, None, None)
    req_o = oracle.decode(pc.req, pc.req_off, pc.req_flags, None, None)
    assert req_g.diff(req_o) == []
    layout = zk.XidTable(np.zeros(n, np.int32), np.full(n, pc.frames, np.uint32))
    xt = ctx.xidtab_build_host(req_g, layout)
    assert np.array_equal(xt.ops, pc.ops)
    rsp_g = ctx.decode_host(pc.rsp, pc.rsp_off, pc.rsp_flags, None, xt)
    rsp_o = oracle.decode(pc.rsp, pc.rsp_off, pc.rsp_flags, None, xt)
    assert rsp_g.diff(rsp_o) == []


@pytest.mark.parametrize("seed", range(6))
def test_many_small_sessions_and_handshake_depths(ctx, oracle, synth, seed):
    """Hundreds of sessions per 8 KiB chunk, empty ones included, with 0..3 (and 'always')
    handshake frames: exercises session lookup, first-frame bookkeeping and per-session results
    inside the decode kernel."""
    import random
    rng = random.Random(seed)
    c = synth.Corpus()
    for _ in range(1500):
        server = rng.random() < 0.5
        hs = rng.choice([0, 1, 1, 2, 3, 0xFFFFFFFF])
        nf = rng.choice([0, 0, 1, 1, 2, 3, 5])
        parts = []
        xm = {}
        for i in range(nf):
            if server:
                pl = synth.connect_request(rng) if rng.random() < 0.3 else (
                    synth.be32(i) + synth.be32(4) + synth.request_body(rng, 4))
            else:
                if rng.random() < 0.3:
                    pl = synth.connect_response(rng)
                else:
                    pl = synth.resp_header(i, i, 0) + synth.response_body(rng, 3)
                    xm[i] = 3
            parts.append(synth.frame(pl))
        tail = rng.choice([b"", b"", b"\x00", b"\x00\x00\x00\x10\x00", b"\xff\xff\xff\xff"])
        c.add(b"".join(parts) + tail, server=server, hs=hs, xidmap=xm)
    g, o = both(ctx, oracle, c)
    assert g.diff(o) == []


def test_sessions_aligned_to_chunk_and_subrange_edges(ctx, oracle, synth):
    """Session and frame boundaries exactly on 256-byte, 8 KiB and batch edges."""
    import random
    rng = random.Random(3)
    c = synth.Corpus()

    def filler(nbytes, i0=0):
        """frames that add up to exactly nbytes (GET_DATA replies with sized payloads)"""
        out, xm, i = [], {}, i0
        left = nbytes
        while left > 0:
            if left < 200:
                size = left
            else:
                size = rng.choice([96, 128, 256, 100, 180])
                if left - size < 96:
                    size = left
            body = size - 4 - 16 - 4 - 68
            assert body >= 0
            out.append(synth.frame(synth.resp_header(i, i, 0) + synth.jbuf(bytes(body)) + synth.stat_bytes(rng)))
            xm[i] = 4
            i += 1
            left -= size
        return b"".join(out), xm

    for nbytes in (8192, 256, 8192 - 256, 16384, 512, 8192 + 256):
        data, xm = filler(nbytes)
        assert len(data) == nbytes
        c.add(data, hs=0, xidmap=xm)
        c.add(b"", hs=1)
    g, o = both(ctx, oracle, c)
    assert g.diff(o) == []
    assert (g.status == 0).all()


def test_cfg3_full_size_one_giant_session(ctx, oracle, synth, zk):
    """Config 3 at BASELINE size: one 1.1 GB session of 32 M tiny frames (20-44 bytes), so the
    splitter has nothing but guesses to go on.  Known structure at full size, the oracle on a
    100 MB prefix."""
    n = 32_000_000
    data = synth.native_cfg3(7, n)
    assert data.size > 1_000_000_000
    cols = ctx.decode_host(data, None, None, None, None)
    assert cols.n_frames == n + 1 and (cols.status == 0).all()
    assert int(cols.sess_consumed[0]) == data.size and int(cols.sess_bad_length_at[0]) == zk._abi.NO_OFFSET
    op = cols.opcode[1:]
    note = op == 0
    assert ((op == 0) | (op == 11)).all() and 0.66 < note.mean() < 0.67
    assert (cols.xid[1:][note] == -1).all() and (cols.zxid[1:][note] == -1).all() and (cols.xid[1:][~note] == -2).all()
    assert cols.n_buf == int(note.sum()) + 1 and cols.n_stat == 0 and cols.n_acl == 0
    assert np.array_equal(cols.buf_frame[1:], np.nonzero(note)[0].astype(np.uint32) + 1)   # one path per notification
    assert (cols.frame_len[1:][~note] == 16).all()
    assert np.array_equal(cols.frame_len[1:][note], 28 + cols.buf_len[1:])
    assert (cols.buf_len[1:] >= 4).all() and (cols.buf_len[1:] <= 12).all()
    assert np.array_equal(cols.buf_off[1:], cols.frame_off[1:][note] + 32)
    assert np.array_equal(np.diff(cols.frame_off.astype(np.int64)), cols.frame_len[:-1].astype(np.int64) + 4)
    assert ((cols.a0[1:][note] >= 1) & (cols.a0[1:][note] <= 4) & (cols.a1[1:][note] == 3)).all()
    # the session folds at this size
    st = ctx.session_stats_host(cols)
    assert int(st.n_replies[0]) == int((~note).sum())
    assert int(st.max_zxid[0]) == int(cols.zxid[1:][~note].max())
    assert [int(x) for x in st.notifications[0]] == [0] + [int((cols.a0[1:][note] == t).sum()) for t in (1, 2, 3, 4)] + [0]
    # full parity on a prefix
    cut = int(cols.frame_off[3_000_001])
    g = ctx.decode_host(data[:cut], None, None, None, None)
    o = oracle.decode(data[:cut], None, None, None, None)
    assert g.diff(o) == [] and g.n_frames == 3_000_001


def test_cfg4_full_size_large_znodes(ctx, oracle, synth, zk):
    """Config 4 at BASELINE size: 1 000 000-byte payloads and 10^4-10^5-child vectors, 8 GB.
    One ~40 MB session is checked against the oracle; ~210 copies of it side by side must each
    decode to the same rows."""
    c = synth.cfg4_large(40, seed=44)
    one = np.frombuffer(c.data, dtype=np.uint8)
    xt1 = c.xidtable()
    g1 = ctx.decode_host(one, None, None, None, xt1)
    assert g1.diff(oracle.decode(one, None, None, None, xt1)) == []
    assert (g1.status == 0).all() and g1.n_frames == 41
    reps = 8_000_000_000 // one.size + 1
    data = np.tile(one, reps)
    assert data.size > 8_000_000_000
    off = np.arange(reps + 1, dtype=np.uint64) * np.uint64(one.size)
    xt = zk.XidTable.from_maps([c.xidmaps[0]] * reps)
    cols = ctx.decode_host(data, off, None, None, xt, copy=False)
    F, B, S = g1.n_frames, g1.n_buf, g1.n_stat
    assert cols.n_frames == reps * F and cols.n_buf == reps * B and cols.n_stat == reps * S
    assert (cols.status == 0).all()
    base = off[:-1].astype(np.int64)
    for name in ("frame_len", "meta", "xid", "zxid", "err", "a0", "a1"):
        assert (getattr(cols, name).reshape(reps, F) == getattr(g1, name)[None, :]).all(), name
    assert (cols.frame_off.astype(np.int64).reshape(reps, F) - base[:, None] == g1.frame_off.astype(np.int64)[None, :]).all()
    assert (cols.session.reshape(reps, F) == np.arange(reps, dtype=np.uint32)[:, None]).all()
    assert (cols.buf_len.reshape(reps, B) == g1.buf_len[None, :]).all()
    assert (cols.buf_off.astype(np.int64).reshape(reps, B) - base[:, None] == g1.buf_off.astype(np.int64)[None, :]).all()
    assert (cols.buf_frame.astype(np.int64).reshape(reps, B) - (np.arange(reps) * F)[:, None] == g1.buf_frame[None, :]).all()
    for name in ("czxid", "mzxid", "ctime", "mtime", "version", "dataLength", "numChildren", "pzxid"):
        assert (getattr(cols, name).reshape(reps, S) == getattr(g1, name)[None, :]).all(), name
    assert (cols.sess_consumed.astype(np.int64) == base + one.size).all()
    assert np.array_equal(cols.sess_first_frame, np.arange(reps + 1, dtype=np.uint32) * F)


def test_half_a_million_tiny_sessions(ctx, oracle, synth):
    """600 000 sessions of 0-3 frames (hundreds of session starts per 8 KiB chunk, a third of the
    sessions empty, some cut mid-frame): session lookup beyond the 32 boundaries a warp keeps in
    registers, first_frame / consumed bookkeeping for sessions that start and end inside one lane's
    sub-range."""
    import random
    rng = random.Random(77)
    ping = synth.frame(synth.resp_header(-2, 9, 0))
    note = synth.frame(synth.resp_header(-1, -1, 0) + synth.be32(3) + synth.be32(3) + synth.jbuf(b"/abc"))
    conn = synth.frame(synth.connect_response(rng))
    kinds = [b"", b"", conn, conn + ping, conn + ping + note, conn + note[:7], ping + note, conn + ping + b"\x7f\xff\xff\xff"]
    hs_of = [1, 1, 1, 1, 1, 1, 0, 1]
    pick = np.frombuffer(bytes(rng.randrange(len(kinds)) for _ in range(600_000)), dtype=np.uint8)
    lens = np.array([len(k) for k in kinds], dtype=np.uint64)[pick]
    off = np.zeros(len(pick) + 1, dtype=np.uint64)
    np.cumsum(lens, out=off[1:])
    data = np.frombuffer(b"".join(kinds[i] for i in pick), dtype=np.uint8)
    hs = np.array(hs_of, dtype=np.uint32)[pick]
    g = ctx.decode_host(data, off, None, hs, None)
    o = oracle.decode(data, off, None, hs, None)
    assert g.diff(o) == []
    assert g.n_frames > 700_000


def test_maximum_frame_length(ctx, oracle, synth, zk):
    """L == 16 MiB is the largest legal length (zk-streams.js:23,49: `len > 16 MiB` is the error);
    one byte more is BAD_LENGTH and ends the session.  A 16 MiB GET_DATA reply spans 2049 chunks."""
    import random
    rng = random.Random(5)
    M = zk._abi.MAX_PACKET
    body_len = M - 16 - 4 - 68
    big = synth.resp_header(1, 7, 0) + synth.jbuf(bytes(body_len)) + synth.stat_bytes(rng, body_len, 0)
    assert len(big) == M
    c = synth.Corpus()
    c.add(synth.frame(synth.connect_response(rng)) + synth.frame(big) + synth.frame(synth.resp_header(-2, 8, 0)),
          xidmap={1: 4})
    c.add(synth.frame(synth.connect_response(rng)) + (M + 1).to_bytes(4, "big") + bytes(100), xidmap={})
    c.add(synth.frame(synth.connect_response(rng)) + synth.frame(big)[:-1], xidmap={1: 4})     # one byte short: stays buffered
    g, o = both(ctx, oracle, c)
    assert g.diff(o) == []
    assert g.n_frames == 5 and list(g.frame_len[:3]) == [len(synth.connect_response(rng)), M, 16]
    assert int(g.buf_len[1]) == body_len and int(g.dataLength[0]) == body_len
    assert int(g.sess_bad_length_at[1]) == int(g.sess_consumed[1]) and int(g.sess_bad_length_at[0]) == zk._abi.NO_OFFSET
    assert int(g.sess_consumed[2]) == int(c.sess_off[2]) + 4 + len(synth.connect_response(rng))


def test_row_prediction_and_its_fallback(ctx, oracle, synth, zk, monkeypatch):
    """Side rows are placed where the splitter predicted; the kernel checks the prediction and
    the pass is repeated with the look-back when it fails.  Well-formed traffic must never need
    the second pass, damaged traffic must get it, and both ways must give the oracle's rows."""
    pc = synth.native_cfg2(17, 24, 4000)
    n = pc.nsessions
    before = ctx.row_mispredicts
    req = ctx.decode_host(pc.req, pc.req_off, pc.req_flags, None, None)
    xt = ctx.xidtab_build_host(req, zk.XidTable(np.zeros(n, np.int32), np.full(n, pc.frames, np.uint32)))
    rsp = ctx.decode_host(pc.rsp, pc.rsp_off, pc.rsp_flags, None, xt)
    c3 = synth.cfg3_tiny(20, 500, seed=8)
    g3, o3 = both(ctx, oracle, c3)
    assert g3.diff(o3) == []
    c4 = synth.cfg4_large(4, payload=200000, children=(2000, 5000))
    g4, o4 = both(ctx, oracle, c4)
    assert g4.diff(o4) == []
    assert ctx.row_mispredicts == before, "well-formed traffic took the fallback"
    assert req.diff(oracle.decode(pc.req, pc.req_off, pc.req_flags, None, None)) == []
    assert rsp.diff(oracle.decode(pc.rsp, pc.rsp_off, pc.rsp_flags, None, xt)) == []
    # damaged frames: prediction must fail somewhere, result still exact
    f = synth.fuzz_corpus(4242, nsessions=200, max_frames=60)
    g, o = both(ctx, oracle, f)
    assert g.diff(o) == []
    assert ctx.row_mispredicts > before
    # handshake depth the walk cannot see (3 connect frames, the third in a later granule)
    import random
    rng = random.Random(1)
    c = synth.Corpus()
    filler = synth.frame(synth.connect_response(rng, passwd_len=9000))
    c.add(filler + filler + synth.frame(synth.connect_response(rng)) + synth.frame(synth.resp_header(-2, 1, 0)), hs=3)
    g, o = both(ctx, oracle, c)
    assert g.diff(o) == [] and g.n_buf == 3
    # and with the prediction switched off the same columns come out
    monkeypatch.setenv("ZKBATCH_NO_PREDICT", "1")
    ctx2 = zk.Context(0)
    try:
        d, off, fl, hs = c3.arrays()
        assert ctx2.decode_host(d, off, fl, hs, c3.xidtable()).diff(o3) == []
        assert ctx2.decode_host(pc.rsp, pc.rsp_off, pc.rsp_flags, None, xt).diff(rsp) == []
        assert ctx2.row_mispredicts == 0
    finally:
        ctx2.close()


@pytest.mark.parametrize("where", [0.0, 0.37, 0.999])
def test_partial_redo_after_a_mispredicted_chunk(ctx, oracle, synth, zk, where):
    """One damaged reply (its child count lies) somewhere in ~60 MB of clean traffic: the rows
    before it stay where the first pass put them, the rest is redone with the look-back."""
    pc = synth.native_cfg2(23, 8, 20000)
    n = pc.nsessions
    xt = pc.xidtable()
    cols = oracle.decode(pc.rsp, pc.rsp_off, pc.rsp_flags, None, xt)
    gc2 = np.nonzero((cols.opcode == 12) & (cols.err == 0) & (cols.a0 > 2))[0]
    f = int(gc2[min(len(gc2) - 1, int(where * len(gc2)))])
    data = pc.rsp.copy()
    pos = int(cols.frame_off[f]) + 4 + 16                      # the child count of that GET_CHILDREN2 reply
    data[pos:pos + 4] = np.frombuffer((int(cols.a0[f]) + 1).to_bytes(4, "big"), dtype=np.uint8)
    before = ctx.row_mispredicts
    g = ctx.decode_host(data, pc.rsp_off, pc.rsp_flags, None, xt)
    o = oracle.decode(data, pc.rsp_off, pc.rsp_flags, None, xt)
    assert g.diff(o) == []
    assert ctx.row_mispredicts == before + 1
    assert int(o.status[f]) == 1                               # the damaged frame is BAD_DECODE, everything else as before
    assert ctx.last_slow_frames >= 1


def test_long_children_vectors_adversarial(ctx, oracle, synth):
    """GET_CHILDREN / GET_CHILDREN2 replies longer than the decode tile: their children are found by
    lanes that guess element boundaries inside 280-byte sub-ranges and repair each other from the
    left.  Names built to mislead the guess (runs of zero bytes, empty names, names longer than 255
    bytes, one name longer than a whole tile, byte patterns that look like prefixes), vectors that end
    right at tile and sub-range boundaries, a count that lies in either direction, and ordinary frames
    before and after - every row must still be the oracle's."""
    import random
    import struct
    rng = random.Random(99)

    def kids_frame(xid, names, op, count=None, stat=True):
        body = synth.resp_header(xid, 7 + xid, 0) + synth.be32(len(names) if count is None else count)
        body += b"".join(struct.pack(">i", len(n) if n or rng.random() < 0.5 else -1) + n for n in names)
        if stat:
            body += synth.stat_bytes(rng, 0, len(names))
        return synth.frame(body)

    def names_plain(n):
        return [b"n%07d" % k + bytes(rng.randrange(0, 16)) for k in range(n)]

    def names_nasty(n):
        out = []
        for k in range(n):
            r = rng.random()
            if r < 0.15:
                out.append(b"")                                            # empty name
            elif r < 0.3:
                out.append(bytes(rng.randrange(1, 40)))                    # nothing but zero bytes
            elif r < 0.4:
                out.append(b"\x00\x00\x00\x05abcde\x00\x00\x00\x03xyz")    # looks like two more elements
            elif r < 0.45:
                out.append(bytes(rng.getrandbits(8) for _ in range(rng.randrange(256, 700))))   # prefix > 255
            else:
                out.append(bytes(rng.choice(b"abcdefghij0123456789") for _ in range(rng.randrange(1, 30))))
        return out

    c = synth.Corpus()
    parts, xm, xid = [synth.frame(synth.connect_response(rng))], {}, 0

    def add(fr, op):
        nonlocal xid
        parts.append(fr)
        xm[xid] = op
        xid += 1

    add(kids_frame(xid, names_plain(3000), 12), 12)
    add(synth.frame(synth.resp_header(xid, 1, 0) + synth.stat_bytes(rng)), 3)              # EXISTS reply in between
    add(kids_frame(xid, names_nasty(4000), 12), 12)
    add(kids_frame(xid, names_nasty(2500), 8, stat=False), 8)                              # GET_CHILDREN: no Stat
    big = [b"a" * 20] * 50 + [bytes(rng.choice(b"xyz") for _ in range(20000))] + [b"b" * 9] * 900   # one name > a tile
    add(kids_frame(xid, big, 12), 12)
    add(kids_frame(xid, names_plain(2000), 12, count=2001), 12)                            # count lies: BAD_DECODE
    add(kids_frame(xid, names_plain(2000), 12, count=1990), 12)                            # fewer than present: Stat read early
    add(kids_frame(xid, names_plain(1500), 12, count=-5), 12)                              # negative count: empty vector
    for n in (697, 698, 699, 700, 701, 1396, 1397):                                        # ends around tile / sub-range edges
        add(kids_frame(xid, [b"0123456789ab"] * n, 12), 12)
    add(synth.frame(synth.resp_header(-2, 3, 0)), 11)
    add(kids_frame(xid, names_nasty(3000), 12), 12)                                        # last frame of the session
    c.add(b"".join(parts), xidmap=xm)
    # the same shapes as the very last bytes of the batch, and in a session of their own without handshake
    c.add(kids_frame(0, names_nasty(1200), 12), hs=0, xidmap={0: 12})
    g, o = both(ctx, oracle, c)
    assert g.diff(o) == []
    assert (o.status == 1).sum() >= 1 and g.n_buf > 15000


def test_crowded_chunks(ctx, oracle, synth):
    """More frames in one 8 KiB chunk than the decode kernel's frame list holds (400), and more
    vector frames than its work list (64): thousands of zero-length frames (legal length, BAD_DECODE
    each), 8-byte frames, and hundreds of tiny GET_CHILDREN / GET_CHILDREN2 / GET_ACL replies with
    0-2 elements back to back, in both directions, with ordinary frames mixed in."""
    import random
    rng = random.Random(7)
    c = synth.Corpus()
    # replies: runs of empty frames, then tiny vector replies
    parts, xm, xid = [synth.frame(synth.connect_response(rng))], {}, 0
    parts += [synth.frame(b"")] * 3000
    for _ in range(2500):
        op = rng.choice([8, 12, 6, 8, 12])
        n = rng.randrange(0, 3)
        if op == 6:
            body = synth.acl_bytes([(rng.randrange(32), b"s", b"i%d" % k) for k in range(n)]) + synth.stat_bytes(rng)
        else:
            body = synth.be32(n) + b"".join(synth.jbuf(b"c%d" % k) for k in range(n))
            if op == 12:
                body += synth.stat_bytes(rng, 0, n)
        parts.append(synth.frame(synth.resp_header(xid, xid + 5, 0) + body))
        xm[xid] = op
        xid += 1
        if rng.random() < 0.3:
            parts.append(synth.frame(synth.resp_header(-2, 1, 0)))
        if rng.random() < 0.05:
            parts += [synth.frame(b"")] * rng.randrange(1, 700)
    c.add(b"".join(parts), xidmap=xm)
    # requests: 8-byte frames (xid + PING / CLOSE_SESSION: 'Unsupported opcode'), CREATEs with 0-3 ACL entries
    parts = [synth.frame(synth.connect_request(rng))]
    for i in range(4000):
        r = rng.random()
        if r < 0.7:
            parts.append(synth.frame(synth.be32(-2) + synth.be32(rng.choice([11, -11]))))
        else:
            ents = [(rng.randrange(32), b"world", b"anyone")] * rng.randrange(0, 4)
            parts.append(synth.frame(synth.be32(i) + synth.be32(1) + synth.jbuf(b"/p") + synth.jbuf(b"") +
                                     synth.acl_bytes(ents) + synth.be32(rng.randrange(4))))
    c.add(b"".join(parts), server=True)
    g, o = both(ctx, oracle, c)
    assert g.diff(o) == []
    assert g.n_frames > 12000 and (g.status == 1).sum() > 5000
