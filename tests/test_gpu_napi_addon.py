"""The N-API addon (node-zkstream_b200/js/zkbatch_napi.c) EXECUTED against the real libzkbatch.so.

There is no node on the box, so tests/napi_host.c stands in for the ~20 N-API functions the addon
uses (values are tagged structs).  The addon object, the host and the codec are linked into one
shared library here; the test then does what js/batch-streams.js does — createContext, decodeHost,
encodeColumns, sessionStats — and compares every typed array the addon hands back with the CPU
oracle.  What the reference side binds in its place: lib/connection-fsm.js:86-91.
"""
import ctypes as C
import importlib
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
PKG = os.path.join(ROOT, "node-zkstream_b200")

# napi_typedarray_type
I8, U8, I32, U32, F64 = 0, 1, 5, 6, 8
NH_NUMBER, NH_OBJECT, NH_TYPEDARRAY, NH_BUFFER, NH_FUNCTION, NH_EXTERNAL = 2, 3, 5, 6, 7, 8


@pytest.fixture(scope="module")
def nh():
    out = os.path.join(HERE, "_napi_host.so")
    srcs = [os.path.join(HERE, "napi_host.c"), os.path.join(PKG, "js", "zkbatch_napi.c")]
    if not os.path.exists(out) or any(os.path.getmtime(s) > os.path.getmtime(out) for s in srcs):
        subprocess.run(["gcc", "-std=gnu99", "-O1", "-Wall", "-fPIC", "-shared", "-o", out] + srcs +
                       ["-I", os.path.join(ROOT, "include"), "-L", PKG, "-lzkbatch", "-Wl,-rpath," + PKG], check=True)
    lib = C.CDLL(out)
    P = C.c_void_p
    for name, res, args in [("nh_exports", P, []), ("nh_module_name", C.c_char_p, []), ("nh_number", P, [C.c_double]),
                            ("nh_null", P, []), ("nh_object", P, []), ("nh_buffer", P, [P, C.c_size_t]),
                            ("nh_typedarray", P, [C.c_int, P, C.c_size_t]), ("nh_set", C.c_int, [P, C.c_char_p, P]),
                            ("nh_get", P, [P, C.c_char_p]), ("nh_kind", C.c_int, [P]), ("nh_num", C.c_double, [P]),
                            ("nh_data", P, [P]), ("nh_len", C.c_size_t, [P]), ("nh_array_type", C.c_int, [P]),
                            ("nh_call", P, [C.c_char_p, C.c_size_t, C.POINTER(P)]), ("nh_last_error", C.c_char_p, []),
                            ("nh_reset", None, [])]:
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    yield lib
    lib.nh_reset()           # finalises the context external: zkb_ctx_destroy


def call(nh, name, *args):
    argv = (C.c_void_p * max(len(args), 1))(*args)
    return nh.nh_call(name.encode(), len(args), argv)


def arr(nh, obj, name, dtype):
    """A typed array property of an addon result as numpy (copy)."""
    v = nh.nh_get(obj, name.encode())
    assert nh.nh_kind(v) == NH_TYPEDARRAY, name
    n = nh.nh_len(v)
    want_type = {np.int32: I32, np.uint32: U32, np.int64: I32, np.uint64: I32, np.uint8: U8}[dtype]
    assert nh.nh_array_type(v) == want_type, name
    if dtype in (np.int64, np.uint64):
        assert n % 2 == 0
        n //= 2           # [lo, hi] pairs of one little-endian 64-bit value
    if n == 0:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(C.cast(nh.nh_data(v), C.POINTER(C.c_uint8)), shape=(n * np.dtype(dtype).itemsize,)).view(dtype).copy()


_DT = {"u1": np.uint8, "i4": np.int32, "u4": np.uint32, "i8": np.int64, "u8": np.uint64}


def test_addon_registers_and_refuses_a_missing_device(nh):
    assert nh.nh_module_name() == b"zkbatch_napi"
    ex = nh.nh_exports()
    for f in ("createContext", "decodeHost", "encodeColumns", "sessionStats"):
        assert nh.nh_kind(nh.nh_get(ex, f.encode())) == NH_FUNCTION
    assert call(nh, "createContext", nh.nh_number(4096)) is None
    assert b"no usable CUDA device" in nh.nh_last_error()


def test_decode_host_through_the_addon_matches_the_oracle(nh, oracle, synth, zk):
    ctx = call(nh, "createContext", nh.nh_number(0))
    assert ctx and nh.nh_kind(ctx) == NH_EXTERNAL
    corpora = [synth.cfg2_mixed(3, 150, seed=31), synth.fuzz_corpus(77, nsessions=40, max_frames=30),
               synth.session_corpus(5, nsessions=10, max_frames=40)]
    checked = 0
    for c in corpora:
        for s, chunk in enumerate(c.chunks):
            data = np.frombuffer(chunk, dtype=np.uint8).copy() if chunk else np.zeros(0, np.uint8)
            server, hs, xm = bool(c.flags[s] & 1), c.hs[s], c.xidmaps[s]
            ops = None
            if xm and not server and max(xm) >= 0:
                ops = np.full(max(xm) + 1, zk._abi.OP_NONE, np.int8)
                for k, v in xm.items():
                    if k >= 0:
                        ops[k] = v
            xt = zk.XidTable([0], [len(ops)], ops) if ops is not None else None
            want = oracle.decode(data, [0, data.size], [1 if server else 0], [hs], xt)
            keep = np.zeros(8, np.uint8) if data.size == 0 else data
            got = call(nh, "decodeHost", ctx, nh.nh_buffer(keep.ctypes.data, data.size), nh.nh_number(1 if server else 0),
                       nh.nh_number(hs), nh.nh_typedarray(I8, ops.ctypes.data, ops.size) if ops is not None else nh.nh_null())
            assert got, nh.nh_last_error()
            for cnt in ("n_frames", "n_stat", "n_buf", "n_acl"):
                assert int(nh.nh_num(nh.nh_get(got, cnt.encode()))) == getattr(want, cnt), (cnt, s)
            assert int(nh.nh_num(nh.nh_get(got, b"consumed"))) == int(want.sess_consumed[0])
            for name, dt in (zk._abi.FRAME_COLS + zk._abi.STAT_COLS + zk._abi.BUF_COLS + zk._abi.ACL_COLS + zk._abi.SESS_COLS):
                a = arr(nh, got, name, _DT[dt])
                assert np.array_equal(a, getattr(want, name)), (name, s)
            checked += want.n_frames
    assert checked > 1000


def test_encode_columns_through_the_addon_matches_the_oracle(nh, oracle, zk):
    ctx = call(nh, "createContext", nh.nh_number(0))
    acl = [dict(perms=["READ", "WRITE"], id=dict(scheme="world", id="anyone")), dict(perms=["ADMIN"], id=dict(scheme="digest", id="u:p"))]
    pkts = []
    for i in range(300):
        pkts.append(dict(xid=i * 3, opcode="SET_DATA", path="/n/%d" % i, data=bytes(range(i % 97)), version=i - 1))
        pkts.append(dict(xid=i * 3 + 1, opcode="GET_CHILDREN2", path="/n/%d" % i, watch=bool(i & 1)))
        pkts.append(dict(xid=i * 3 + 2, opcode="CREATE", path="/c/%d" % i, data=b"x" * (i % 5), acl=acl, flags=["EPHEMERAL"] if i & 1 else []))
    pkts += [dict(xid=-2, opcode="PING"), dict(xid=7, opcode="CLOSE_SESSION"),
             dict(xid=-8, opcode="SET_WATCHES", relZxid=bytes([0, 0, 0, 0, 0, 0, 5, 23]),
                  events=dict(dataChanged=["/a"], createdOrDestroyed=[], childrenChanged=["/b", "/cc"]))]
    rb = zk.RequestBatch.from_packets(pkts)
    od, oo, os_ = oracle.encode(rb)
    cols = nh.nh_object()
    keep = []
    for name, dt in zk._abi.REQ_COLS + zk._abi.ELEM_COLS + zk._abi.REQ_ACL_COLS:
        a = np.ascontiguousarray(rb.cols[name], dtype=_DT.get(dt, np.int8))
        if a.size == 0:
            a = np.zeros(1, a.dtype)[:0]
        keep.append(a)
        n = a.size * (2 if a.dtype.itemsize == 8 else 1)
        t = I8 if a.dtype.itemsize == 1 else (I32 if a.dtype != np.uint32 else U32)
        nh.nh_set(cols, name.encode(), nh.nh_typedarray(t, a.ctypes.data if a.size else None, n))
    arena = np.ascontiguousarray(rb.arena)
    nh.nh_set(cols, b"arena", nh.nh_buffer(arena.ctypes.data, arena.size))
    got = call(nh, "encodeColumns", ctx, cols)
    assert got, nh.nh_last_error()
    b = nh.nh_get(got, b"bytes")
    assert nh.nh_kind(b) == NH_BUFFER and nh.nh_len(b) == od.size
    gb = np.ctypeslib.as_array(C.cast(nh.nh_data(b), C.POINTER(C.c_uint8)), shape=(od.size,))
    assert np.array_equal(gb, od)
    fo = nh.nh_get(got, b"frame_off")
    assert nh.nh_array_type(fo) == F64 and nh.nh_len(fo) == rb.n + 1
    gfo = np.ctypeslib.as_array(C.cast(nh.nh_data(fo), C.POINTER(C.c_double)), shape=(rb.n + 1,))
    assert np.array_equal(gfo.astype(np.uint64), oo)
    assert np.array_equal(arr(nh, got, "status", np.uint8), os_)


def test_session_stats_through_the_addon_matches_the_oracle(nh, oracle, synth, zk):
    ctx = call(nh, "createContext", nh.nh_number(0))
    c = synth.session_corpus(9, nsessions=6, max_frames=80)
    for s, chunk in enumerate(c.chunks):
        data = np.frombuffer(chunk, dtype=np.uint8).copy() if chunk else np.zeros(8, np.uint8)[:0]
        xm = c.xidmaps[s]
        ops = np.full((max(xm) + 1) if xm else 1, zk._abi.OP_NONE, np.int8)
        for k, v in xm.items():
            ops[k] = v
        xt = zk.XidTable([0], [len(ops)], ops)
        want_cols = oracle.decode(data, [0, data.size], [0], [c.hs[s]], xt)
        keep = np.zeros(8, np.uint8) if data.size == 0 else data
        got = call(nh, "decodeHost", ctx, nh.nh_buffer(keep.ctypes.data, data.size), nh.nh_number(0), nh.nh_number(c.hs[s]),
                   nh.nh_typedarray(I8, ops.ctypes.data, ops.size))
        assert got, nh.nh_last_error()
        for flags in (0, zk._abi.STATS_UNTIL_ERROR):
            st = call(nh, "sessionStats", ctx, got, nh.nh_number(1), nh.nh_number(flags))
            assert st, nh.nh_last_error()
            want = oracle.session_stats(want_cols, 1, flags)
            assert np.array_equal(arr(nh, st, "max_zxid", np.int64), want.max_zxid)
            assert np.array_equal(arr(nh, st, "n_replies", np.uint32), want.n_replies)
            assert np.array_equal(arr(nh, st, "notifications", np.uint32), want.notifications.reshape(-1))
