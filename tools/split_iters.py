"""Development aid: dirty granules after each scan of the last split, per direction of a cfg2 corpus.

    [ZKBATCH_SPLIT_RUN=n] python tools/split_iters.py [gb]
"""
import ctypes as C
import importlib
import sys

sys.path.insert(0, ".")
zk = importlib.import_module("node-zkstream_b200")
synth = importlib.import_module("node-zkstream_b200.synth")
dev = importlib.import_module("node-zkstream_b200.device")

gb = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
S = max(16, int(gb * 1e9 / 7.79e6))
pc = synth.native_cfg2(1, S, 20000)
ctx = zk.Context(0)
lib = ctx._lib
lib.zkb_debug_split_iters.argtypes = [C.c_void_p, C.POINTER(C.c_uint * 64)]
device = "cuda:0"
for label, data, off, fl, xt in (("request streams", pc.req, pc.req_off, pc.req_flags, None),
                                 ("response streams", pc.rsp, pc.rsp_off, pc.rsp_flags, pc.xidtable())):
    d = dev.to_device(data, device, pad=64)
    ds = dev.DeviceSessions(data.size, off, fl, None, device)
    dx = dev.DeviceXidTable(xt, device) if xt is not None else None
    dc = dev.decode_dev_autosize(ctx, d, int(data.size), ds, dx, device=device)
    dev.decode_dev(ctx, d, int(data.size), ds, dx, dc)
    it = (C.c_uint * 64)()
    lib.zkb_debug_split_iters(ctx._h, C.byref(it))
    print(label, "bytes", data.size, "(n_dirty, n_rewalk) per scan:", [(it[2 * i], it[2 * i + 1]) for i in range(32) if it[2 * i] or i == 0])
