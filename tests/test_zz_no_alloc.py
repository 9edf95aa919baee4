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


def _build_both(ctx, zk, dev, dc, layout, nsessions):
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
                                      None, err, len(err))
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
