"""CPU-only checks: the shared library loads and exports exactly the C ABI that
include/zkbatch.h declares (no compute calls), and the host-side logic around it."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "zkbatch.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(zkb_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(zk):
    names = declared_functions()
    assert len(names) >= 15
    lib = ctypes.CDLL(zk.LIB_PATH)
    for n in names:
        assert getattr(lib, n) is not None, n
    bound = {n for n, _, _ in zk._abi.SYMBOLS}
    assert bound == set(names)


def test_library_exports_nothing_else(zk):
    import subprocess
    out = subprocess.run(["nm", "-D", "--defined-only", zk.LIB_PATH], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert {e for e in exported if not e.startswith("zkb_")} <= {"_init", "_fini"}


def test_no_gpu_means_loud_failure(zk):
    """Without a CUDA device the context cannot be created; there is no CPU path to fall back to."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(zk.ZKBatchError) as e:
        zk.Context(0)
    assert e.value.code == zk._abi.ENODEV


def test_struct_layouts_match_header(zk):
    a = zk._abi
    assert ctypes.sizeof(a.Sessions) == 32
    assert ctypes.sizeof(a.XidTab) == 40
    assert ctypes.sizeof(a.Decoded) == 8 * 8 + 8 * (9 + 12 + 3 + 2 + 3 + 1)      # + buf_rel
    assert ctypes.sizeof(a.Requests) == 8 * 26


def test_request_batch_validation(zk):
    from importlib import import_module
    enc = import_module("node-zkstream_b200.encode")
    RB = zk.RequestBatch
    with pytest.raises(enc.EncodeAssertionError):
        RB.from_packets([dict(opcode="GET_DATA", path="/a", watch=False)])            # xid missing
    with pytest.raises(enc.EncodeAssertionError):
        RB.from_packets([dict(xid=1, opcode="GET_DATA", path="/a", watch=1)])          # watch not bool
    with pytest.raises(enc.EncodeAssertionError):
        RB.from_packets([dict(xid=1, opcode="CREATE", path="/a", data=b"", acl=[], flags=["BOGUS"])])
    with pytest.raises(enc.EncodeAssertionError):
        RB.from_packets([dict(xid=1, opcode="SET_DATA", path="/a", data="str", version=0)])
    b = RB.from_packets([dict(xid=1, opcode="CREATE", path="/a", data=b"xy", flags=["EPHEMERAL"],
                              acl=[dict(perms=["read", "ADMIN"], id=dict(scheme="world", id="anyone"))])])
    assert b.cols["a0"][0] == 1 and b.cols["acl_perms"][0] == 17 and b.n_acl == 1


def test_write_long_padding_quirk(zk):
    """writeLong left-pads with zeros, it does not sign-extend (lib/jute-buffer.js:162-163)."""
    from importlib import import_module
    enc = import_module("node-zkstream_b200.encode")
    assert enc.long_to_int64(bytes.fromhex("0517")) == 0x517
    assert enc.long_to_int64(-1) == 0xff                    # the reference writes 00..00ff
    assert enc.long_to_int64(0x1234567890) == 0x1234567890
    assert enc.long_to_int64(zk.LongDate(bytes(7) + b"\x09")) == 9
    with pytest.raises(AssertionError):
        enc.long_to_int64(bytes(9))


def test_xid_table_from_maps(zk):
    t = zk.XidTable.from_maps([{0: 1, 2: 4}, {}, {5: 8}])
    assert list(t.base) == [0, 0, 5] and list(t.count) == [3, 0, 1]
    assert list(t.ops) == [1, -128, 4, 8] and list(t.off) == [0, 3, 3]


def test_partition_and_merge(zk, synth, oracle):
    from importlib import import_module
    shard = import_module("node-zkstream_b200.shard")
    c = synth.fuzz_corpus(3, nsessions=30, max_frames=20)
    d, off, fl, hs = c.arrays()
    xt = c.xidtable()
    full = oracle.decode(d, off, fl, hs, xt)
    for world in (1, 2, 3, 8):
        parts = shard.partition_sessions(off, world)
        assert parts[0][0] == 0 and parts[-1][1] == 30
        assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
        cols, bb, sb = [], [], []
        for s0, s1 in parts:
            sd, so, sf, sh, sx = shard.shard_inputs(d, off, fl, hs, xt, s0, s1)
            cols.append(oracle.decode(sd, so, sf, sh, sx))
            bb.append(int(off[s0]))
            sb.append(s0)
        assert shard.merge(cols, bb, sb).diff(full) == []


def test_restore_derived_puts_back_what_compact_replay_leaves_out(zk, synth, oracle):
    """Columns.restore_derived: frame_off, session, the *_frame columns and buf_off rebuilt from the
    columns a compact replay result still carries, on corpora with damaged frames, empty sessions,
    handshakes of several depths and every opcode - must equal the full columns."""
    for corpus in (synth.cfg2_mixed(6, 300, seed=5), synth.fuzz_corpus(11, nsessions=120, max_frames=50),
                   synth.cfg3_tiny(12, 80), synth.session_corpus(3, 20, 60)):
        d, off, fl, hs = corpus.arrays()
        full = oracle.decode(d, off, fl, hs, corpus.xidtable())
        c = zk.Columns()
        for k, v in full.__dict__.items():
            setattr(c, k, v.copy() if isinstance(v, np.ndarray) else v)
        c.buf_rel = (full.buf_off - full.frame_off[full.buf_frame]).astype(np.uint32)
        for name in ("frame_off", "session", "stat_frame", "buf_frame", "acl_frame", "buf_off"):
            setattr(c, name, np.zeros(0, getattr(full, name).dtype))
        assert c.restore_derived(off).diff(full) == []
