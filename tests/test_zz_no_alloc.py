"""The allocation-free xid-table build (zkb_xidtab_build_noalloc) against the ordinary
device entry point (zkb_xidtab_build_dev), through the C ABI.

The no-alloc variant is for embedders that cannot let the library call cudaMalloc (their
own pool owns the device memory, or their service loop cannot afford the device-wide stall
an allocation causes): its scratch is a fixed array inside the library image, shared by all
calls.  So the things to check are that the result is identical to the allocating entry
point, that a second call is not confused by what the first one left in the shared array,
and that a table larger than the array is refused rather than overrun.

Runs last (file name) and is self-contained: it passes alone in a fresh process and after
the rest of the suite.
"""
import ctypes as C
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _tables(ctx, zk, synth, dev, seed, nsessions, frames):
    """Decode the request side of a paired corpus on the device; return what both builders need."""
    import torch
    device = "cuda:0"
    pc = synth.native_cfg2(seed, nsessions, frames)
    d_req = dev.to_device(pc.req, device, pad=64)
    ds = dev.DeviceSessions(pc.req.size, pc.req_off, pc.req_flags, None, device)
    dc = dev.decode_dev_autosize(ctx, d_req, int(pc.req.size), ds, None, device=device)
    layout = zk.XidTable(np.zeros(nsessions, np.int32), np.full(nsessions, frames, np.uint32))
    return pc, dc, layout, (d_req, ds)


def _build_both(ctx, zk, dev, dc, layout, nsessions, stream=None):
    import torch
    device = "cuda:0"
    lib = ctx._lib
    # existing entry point (context scratch, obtained with cudaMalloc)
    dx_a = dev.DeviceXidTable(layout, device)
    dx_a.ops.fill_(77)
    dev.xidtab_build_dev(ctx, dc, nsessions, dx_a)
    # new entry point: no context, caller's stream (NULL = default stream), fixed scratch
    dx_b = dev.DeviceXidTable(layout, device)
    dx_b.ops.fill_(55)
    torch.cuda.synchronize()
    err = C.create_string_buffer(256)
    rc = lib.zkb_xidtab_build_noalloc(C.byref(dc.struct), nsessions, C.byref(dx_b.struct), dx_b.ops.data_ptr(),
                                      stream.cuda_stream if stream is not None else None, err, len(err))
    assert rc == 0, err.value
    torch.cuda.synchronize()
    n = layout.nops
    a = dx_a.ops[:n].cpu().numpy().view(np.int8)
    b = dx_b.ops[:n].cpu().numpy().view(np.int8)
    return a, b


def test_noalloc_matches_allocating_entry_point_twice():
    zk = importlib.import_module("node-zkstream_b200")
    synth = importlib.import_module("node-zkstream_b200.synth")
    dev = importlib.import_module("node-zkstream_b200.device")
    ctx = zk.Context(0)
    try:
        # first call: a table that fills a good part of the fixed array
        pc1, dc1, lay1, keep1 = _tables(ctx, zk, synth, dev, 11, 24, 3000)
        a1, b1 = _build_both(ctx, zk, dev, dc1, lay1, 24)
        assert np.array_equal(a1, b1)
        assert np.array_equal(b1, pc1.ops)          # and both equal what the generator wrote
        # second call right after, different and smaller: stale keys of call 1 must not leak in
        pc2, dc2, lay2, keep2 = _tables(ctx, zk, synth, dev, 12, 5, 700)
        a2, b2 = _build_both(ctx, zk, dev, dc2, lay2, 5)
        assert np.array_equal(a2, b2)
        assert np.array_equal(b2, pc2.ops)
        # and the first corpus again
        a3, b3 = _build_both(ctx, zk, dev, dc1, lay1, 24)
        assert np.array_equal(a3, b3) and np.array_equal(b3, b1)
    finally:
        ctx.close()


def test_noalloc_refuses_tables_beyond_the_fixed_scratch():
    zk = importlib.import_module("node-zkstream_b200")
    _abi = zk._abi
    lib = zk.load()
    d = _abi.Decoded()
    t = _abi.XidTab()
    t.nops = _abi.NOALLOC_MAX_XIDS + 1
    err = C.create_string_buffer(256)
    rc = lib.zkb_xidtab_build_noalloc(C.byref(d), 1, C.byref(t), None, None, err, len(err))
    assert rc == _abi.ERANGE
    assert b"exceeds" in err.value


def test_noalloc_on_a_stream_of_the_callers():
    """The documented use: the embedder queues the build on its own (non-default) stream."""
    import torch
    zk = importlib.import_module("node-zkstream_b200")
    synth = importlib.import_module("node-zkstream_b200.synth")
    dev = importlib.import_module("node-zkstream_b200.device")
    ctx = zk.Context(0)
    try:
        st = torch.cuda.Stream(device="cuda:0")
        pc, dc, lay, keep = _tables(ctx, zk, synth, dev, 13, 9, 1500)
        for _ in range(2):
            a, b = _build_both(ctx, zk, dev, dc, lay, 9, stream=st)
            assert np.array_equal(a, b)
            assert np.array_equal(b, pc.ops)
    finally:
        ctx.close()


def test_noalloc_accepts_a_table_of_exactly_the_maximum():
    """nops == ZKB_NOALLOC_MAX_XIDS is the largest table the fixed array holds: accepted, every
    slot used (the last session's xids land in the last slots), same answer as the allocating twin."""
    zk = importlib.import_module("node-zkstream_b200")
    synth = importlib.import_module("node-zkstream_b200.synth")
    dev = importlib.import_module("node-zkstream_b200.device")
    ctx = zk.Context(0)
    try:
        ns, per = 1024, 1024
        assert ns * per == zk._abi.NOALLOC_MAX_XIDS
        pc = synth.native_cfg2(21, ns, 40)
        d_req = dev.to_device(pc.req, "cuda:0", pad=64)
        ds = dev.DeviceSessions(pc.req.size, pc.req_off, pc.req_flags, None, "cuda:0")
        dc = dev.decode_dev_autosize(ctx, d_req, int(pc.req.size), ds, None, device="cuda:0")
        lay = zk.XidTable(np.zeros(ns, np.int32), np.full(ns, per, np.uint32))
        a, b = _build_both(ctx, zk, dev, dc, lay, ns)
        assert np.array_equal(a, b)
        got = b.reshape(ns, per)
        assert np.array_equal(got[:, :40].reshape(-1), pc.ops)
        assert (got[:, 40:] == zk._abi.OP_NONE).all()
    finally:
        ctx.close()


def test_layout_that_points_past_the_table_is_ignored_by_both_builders():
    """off/count live in device memory, so the host cannot validate them: entries that would land
    beyond nops must be dropped by the kernel (they used to be written past the scratch)."""
    import torch
    zk = importlib.import_module("node-zkstream_b200")
    synth = importlib.import_module("node-zkstream_b200.synth")
    dev = importlib.import_module("node-zkstream_b200.device")
    ctx = zk.Context(0)
    try:
        ns, frames = 6, 900
        pc, dc, lay, keep = _tables(ctx, zk, synth, dev, 14, ns, frames)
        good_a, good_b = _build_both(ctx, zk, dev, dc, lay, ns)
        assert np.array_equal(good_a, good_b)
        # same arrays, but the library is told the table is shorter than the layout claims
        short = zk.XidTable(lay.base, lay.count)
        short.nops = (ns - 2) * frames + 17
        for noalloc in (False, True):
            dx = dev.DeviceXidTable(lay, "cuda:0")
            dx.struct.nops = short.nops
            guard = dx.ops.clone()
            guard.fill_(91)
            dx.ops.copy_(guard)
            if noalloc:
                err = C.create_string_buffer(256)
                rc = ctx._lib.zkb_xidtab_build_noalloc(C.byref(dc.struct), ns, C.byref(dx.struct), dx.ops.data_ptr(),
                                                       None, err, len(err))
                assert rc == 0, err.value
            else:
                dev.xidtab_build_dev(ctx, dc, ns, dx)
            torch.cuda.synchronize()
            got = dx.ops.cpu().numpy().view(np.int8)
            assert np.array_equal(got[:short.nops], good_a[:short.nops])
            assert (got[short.nops:lay.nops] == 91).all()      # nothing written past the table
        # and the shared array is still sane for the next ordinary call
        a, b = _build_both(ctx, zk, dev, dc, lay, ns)
        assert np.array_equal(a, good_a) and np.array_equal(b, good_a)
    finally:
        ctx.close()
