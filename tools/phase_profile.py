"""Development aid: per-phase SM cycles of k_decode, from a library built with
-DZK_PROFILE_PHASES (see the comment in csrc/zk_decode.cu).

    ZKBATCH_LIB=tools/libzkbatch_prof.so python tools/phase_profile.py [gb]
"""
import ctypes as C
import importlib
import sys

import numpy as np

sys.path.insert(0, ".")
zk = importlib.import_module("node-zkstream_b200")
synth = importlib.import_module("node-zkstream_b200.synth")
dev = importlib.import_module("node-zkstream_b200.device")

gb = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
S = max(16, int(gb * 1e9 / 7.79e6))
pc = synth.native_cfg2(1, S, 20000)
ctx = zk.Context(0)
lib = ctx._lib
lib.zkb_debug_phase_cycles.argtypes = [C.c_void_p, C.POINTER(C.c_ulonglong * 8)]
device = "cuda:0"
names = ["ticket", "issue+prefetch", "tile wait", "frame loop", "counts", "look-back", "emission", "chunks"]
for label, data, off, fl, xt in (("request streams", pc.req, pc.req_off, pc.req_flags, None),
                                 ("response streams", pc.rsp, pc.rsp_off, pc.rsp_flags, pc.xidtable())):
    d = dev.to_device(data, device, pad=64)
    ds = dev.DeviceSessions(data.size, off, fl, None, device)
    dx = dev.DeviceXidTable(xt, device) if xt is not None else None
    dc = dev.decode_dev_autosize(ctx, d, int(data.size), ds, dx, device=device)
    hist = (C.c_ulonglong * 192)()
    for it in range(3):
        if it == 2:
            lib.zkb_debug_hist(C.byref(hist), 1)
        dev.decode_dev(ctx, d, int(data.size), ds, dx, dc)
    import torch
    torch.cuda.synchronize()
    lib.zkb_debug_hist(C.byref(hist), 0)
    for hi, nm in enumerate(("issue->AGG, fast chunks", "issue->AGG, chunks with generic-parser frames", "look-back wait")):
        h = [hist[hi * 64 + b] for b in range(64)]
        tot = max(1, sum(h))
        cum, qs = 0, {}
        for b, v in enumerate(h):
            cum += v
            for q in (0.5, 0.9, 0.99, 0.999, 1.0):
                if q not in qs and cum >= q * tot:
                    qs[q] = (b + 1) * 2048
        print("   hist %-46s n=%8d  p50 %6d p90 %6d p99 %6d p99.9 %6d max %6d cycles" % (
            nm, sum(h), qs[0.5], qs[0.9], qs[0.99], qs[0.999], qs[1.0]))
    sp = (C.c_ulonglong * 8)()
    lib.zkb_debug_split_cycles.argtypes = [C.c_void_p, C.POINTER(C.c_ulonglong * 8)]
    lib.zkb_debug_split_cycles(ctx._h, C.byref(sp))
    if sp[2]:
        print(label, "k_split<SPEC> per warp: guess %.0f cycles, walk %.0f cycles (%d warps)" % (
            sp[0] / sp[2], sp[1] / sp[2], sp[2]))
        lanes = sp[2] * 32
        print("     per lane: %.2f segments of 128 B scanned, %.2f candidates, %.0f cycles in far validation" % (
            sp[4] / lanes, sp[5] / lanes, sp[3] / lanes))
    it = (C.c_uint * 64)()
    lib.zkb_debug_split_iters.argtypes = [C.c_void_p, C.POINTER(C.c_uint * 64)]
    lib.zkb_debug_split_iters(ctx._h, C.byref(it))
    print(label, "split iterations (n_dirty, n_rewalk):", [(it[2 * i], it[2 * i + 1]) for i in range(32) if it[2 * i] or i == 0])
    out = (C.c_ulonglong * 8)()
    lib.zkb_debug_phase_cycles(ctx._h, C.byref(out))
    n = max(1, out[7])
    tot = sum(out[i] for i in range(7))
    print(label, "chunks", out[7], "avg cycles per chunk %.0f" % (tot / n))
    for i in range(7):
        print("   %-16s %8.0f cycles  %5.1f%%" % (names[i], out[i] / n, 100.0 * out[i] / max(tot, 1)))
