"""Pins the CPU oracle (oracle/zkoracle.c) — and with it everything the GPU tests compare
against — to the reference: its own test vectors, and tests/golden/*.json, which holds what
the reference's unmodified JavaScript produced (generator: oracle/make_golden.py)."""
import base64
import json
import os

import numpy as np
import pytest

from _canon import LongDate, materialised_events, zk

GOLD = os.path.join(os.path.dirname(__file__), "golden")

CAPTURE1 = [  # reference test/streams.test.js:21-27
    ("send", "AAAALQAAAAAAAAAAAAAAAAAAdTAAAAAAAAAAAAAAABAAAAAAAAAAAAAAAAAAAAAAAA=="),
    ("recv", "AAAAJQAAAAAAAHUwAVWjqFbbAAAAAAAQh19uvwgo25o9B6hUkSvqKQA="),
    ("send", "AAAADgAAAAEAAAAIAAAAAS8A"),
    ("recv", "AAAAKAAAAAEAAAAAAAAFFwAAAAAAAAACAAAACXpvb2tlZXBlcgAAAANmb28="),
]
DECODED1 = [  # reference test/streams.test.js:28-56
    {"lastZxidSeen": bytes(8), "passwd": bytes(16), "protocolVersion": 0, "sessionId": bytes(8), "timeOut": 30000},
    {"passwd": base64.b64decode("h19uvwgo25o9B6hUkSvqKQ=="), "protocolVersion": 0,
     "sessionId": base64.b64decode("AVWjqFbbAAA="), "timeOut": 30000},
    {"xid": 1, "opcode": "GET_CHILDREN", "path": "/", "watch": False},
    {"xid": 1, "opcode": "GET_CHILDREN", "err": "OK", "zxid": bytes([0, 0, 0, 0, 0, 0, 5, 0x17]),
     "children": ["zookeeper", "foo"]},
]


def test_capture1_decodes_to_decoded1(oracle, synth):
    """The reference's one offline codec test (test/streams.test.js:58-82)."""
    b = [base64.b64decode(x) for _, x in CAPTURE1]
    c = synth.Corpus()
    c.add(b[0] + b[2], server=True)
    c.add(b[1] + b[3], xidmap={1: 8})
    d, off, fl, hs = c.arrays()
    cols = oracle.decode(d, off, fl, hs, c.xidtable())
    m = zk.Materialiser(cols, c.data)
    got = [p for k, p in m.events() if k == "pkt"]
    assert got == [DECODED1[0], DECODED1[2], DECODED1[1], DECODED1[3]]
    # t.strictEqual(len, zkb.remainder().length): every frame was consumed exactly
    assert list(cols.sess_consumed) == [len(b[0] + b[2]), len(c.data)]


def test_nasty_bad_lengths(oracle, synth):
    """reference test/nasty.test.js:125-126,148,170"""
    c = synth.Corpus()
    c.add(bytes.fromhex("4000") + bytes.fromhex("4000"))
    c.add(bytes.fromhex("000000000102"))
    c.add(bytes.fromhex("fffffffe0102"))
    d, off, fl, hs = c.arrays()
    cols = oracle.decode(d, off, fl, hs, None)
    ev = [[(k, getattr(e, "code", None)) for k, e in zk.Materialiser(cols, c.data).events(s)] for s in range(3)]
    assert ev[0] == [("error", "BAD_LENGTH")]
    assert ev[1] == [("error", "BAD_DECODE")]      # zero length is legal, the decode then throws
    assert ev[2] == [("error", "BAD_LENGTH")]


def test_perms_quirk():
    """cdrwa = 31 reads back as all five names (test/basic.test.js:180-181), any even value as none."""
    from importlib import import_module
    mat = import_module("node-zkstream_b200.materialise")
    assert mat.perms_from_raw(31) == ["READ", "WRITE", "CREATE", "DELETE", "ADMIN"]
    assert mat.perms_from_raw(1) == ["READ", "WRITE", "CREATE", "DELETE", "ADMIN"]
    assert mat.perms_from_raw(0x1e) == []


def _load(name):
    with open(os.path.join(GOLD, name)) as f:
        return json.load(f)


def test_oracle_matches_reference_decode_fixtures(oracle, synth):
    """Every session in ref_decode.json: packets, key order, errors and leftover byte count."""
    g = _load("ref_decode.json")
    assert len(g["sessions"]) > 100
    c = synth.Corpus()
    for s in g["sessions"]:
        c.add(bytes.fromhex(s["hex"]), server=s["isServer"], hs=s["hs"],
              xidmap={int(k): v for k, v in s["xidMap"].items()})
    d, off, fl, hs = c.arrays()
    cols = oracle.decode(d, off, fl, hs, c.xidtable())
    for i, s in enumerate(g["sessions"]):
        assert materialised_events(cols, c.data, i) == s["events"], (s["corpus"], i)
        assert int(off[i + 1]) - int(cols.sess_consumed[i]) == s["used"], (s["corpus"], i)
    # the faithful (chunked, copying) mode of the oracle gives the same columns
    assert cols.diff(oracle.decode(d, off, fl, hs, c.xidtable(), mode=1, chunk=37)) == []


def unjson(v):
    if isinstance(v, dict):
        if "$hex" in v:
            return bytes.fromhex(v["$hex"])
        return {k: unjson(x) for k, x in v.items()}
    if isinstance(v, list):
        return [unjson(x) for x in v]
    return v


def test_oracle_matches_reference_encode_fixtures(oracle):
    g = _load("ref_encode.json")
    for p in g["packets"]:
        pkt = unjson(p["pkt"])
        batch = zk.RequestBatch.from_packets([pkt], p["handshaking"], p["isServer"])
        data, off, st = oracle.encode(batch)
        if "frame" in p["result"]:
            assert st[0] == 0
            assert bytes(data).hex() == p["result"]["frame"], pkt
        else:
            assert st[0] == 1 and off[1] == 0
            assert "Unsupported opcode" in p["result"]["throw"]


def test_survey_known_answers(oracle):
    """Known-answer frames obtained by executing the reference source."""
    kat = [
        (dict(xid=1, opcode="GET_CHILDREN", path="/", watch=False), "0000000e0000000100000008000000012f00"),
        (dict(xid=-2, opcode="PING"), "00000008fffffffe0000000b"),
        (dict(xid=11, opcode="CLOSE_SESSION"), "000000080000000bfffffff5"),
        (dict(xid=5, opcode="EXISTS", path="/foo", watch=True), "000000110000000500000003000000042f666f6f01"),
        (dict(xid=8, opcode="DELETE", path="/foo", version=0), "000000140000000800000002000000042f666f6f00000000"),
        (dict(xid=4, opcode="SET_DATA", path="/foo", data=b"hi", version=-1),
         "0000001a0000000400000005000000042f666f6f000000026869ffffffff"),
        (dict(xid=3, opcode="CREATE", path="/e", data=b"", acl=[dict(id=dict(scheme="world", id="anyone"),
              perms=["read", "write", "create", "delete", "admin"])], flags=["EPHEMERAL", "SEQUENTIAL"]),
         "000000310000000300000001000000022f65ffffffff000000010000001f00000005776f726c6400000006616e796f6e6500000003"),
        (dict(xid=-8, opcode="SET_WATCHES", relZxid=bytes.fromhex("0000000000000517"),
              events=dict(dataChanged=["/a"], createdOrDestroyed=[], childrenChanged=["/b", "/cc"])),
         "0000002ffffffff800000065000000000000051700000001000000022f610000000000000002000000022f62000000032f6363"),
    ]
    batch = zk.RequestBatch.from_packets([k for k, _ in kat])
    data, off, st = oracle.encode(batch)
    for i, (_, want) in enumerate(kat):
        assert bytes(data[int(off[i]):int(off[i + 1])]).hex() == want
    b = zk.RequestBatch.from_packets([dict(protocolVersion=0, lastZxidSeen=bytes(8), timeOut=30000,
                                           sessionId=bytes(8), passwd=bytes(8))], True, False)
    data, off, st = oracle.encode(b)
    assert bytes(data).hex() == "00000024" + "00000000" + "00" * 8 + "00007530" + "00" * 8 + "00000008" + "00" * 8


def test_encode_then_decode_roundtrip(oracle, synth):
    """writeRequest output read back by readRequest gives the packets back (empty buffers come
    back empty although they travel as length -1)."""
    g = _load("ref_encode.json")
    pkts = [unjson(p["pkt"]) for p in g["packets"] if not p["handshaking"] and "frame" in p["result"]]
    batch = zk.RequestBatch.from_packets(pkts)
    data, off, st = oracle.encode(batch)
    cols = oracle.decode(data, [0, len(data)], [1], [0], None)
    got = [p for k, p in zk.Materialiser(cols, bytes(data)).events() if k == "pkt"]
    back = [p for p in pkts if p["opcode"] not in ("PING", "CLOSE_SESSION")]   # readRequest cannot read these
    assert len(got) == len(back)
    for a, b in zip(got, back):
        assert a["xid"] == b["xid"] and a["opcode"] == b["opcode"]
        for k in ("path", "watch", "version", "data"):
            if k in b:
                assert a[k] == b[k]
        if b["opcode"] == "CREATE":
            assert a["flags"] == b["flags"]
            assert [x["id"] for x in a["acl"]] == [x["id"] for x in b["acl"]]


def test_longdate():
    ld = LongDate(bytes.fromhex("0000014b3d2f9c00"))
    assert ld.toDate().year == 2015 and ld.toBuffer() == bytes.fromhex("0000014b3d2f9c00")


def test_whole_batch_digest_matches_the_columns():
    """zko_digest (threaded decode folded into one digest per column: bench.py's parity gate) and the
    tensor-arithmetic digest the bench computes over the GPU's columns (verify.digest_tensor, run
    here on CPU tensors) must both equal the digest of the oracle's own columns."""
    import importlib
    import torch
    import zkoracle
    synth = importlib.import_module("node-zkstream_b200.synth")
    ver = importlib.import_module("node-zkstream_b200.verify")
    for corpus in (synth.cfg2_mixed(5, 120, seed=3), synth.fuzz_corpus(5, nsessions=60, max_frames=30)):
        d, off, fl, hs = corpus.arrays()
        xt = corpus.xidtable()
        cols = zkoracle.decode(d, off, fl, hs, xt)
        want = zkoracle.digest_columns(cols)
        for nt in (1, 4):
            got, cnt = zkoracle.digest(d, off, fl, hs, xt, nt)
            assert got == want
            assert cnt == (cols.n_frames, cols.n_stat, cols.n_buf, cols.n_acl)
        for name in ver.COLUMNS:
            a = getattr(cols, name)
            t = torch.from_numpy(a.view(np.int32 if a.dtype.itemsize == 4 else np.int64).copy())
            assert ver.digest_tensor(t, a.size, chunk=257) == want[name], name
