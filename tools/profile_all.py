"""Development aid: one pass of every kernel family over a cfg2 corpus, for an `ncu` capture.

    python tools/profile_all.py [GB] [passes]

decode(request streams) -> xid table build -> decode(response streams) -> session folds ->
re-encode of the decoded requests.  Prints the CUDA-event time of each stage of the last pass.
"""
import importlib
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
zk = importlib.import_module("node-zkstream_b200")
dev = importlib.import_module("node-zkstream_b200.device")
wls = importlib.import_module("node-zkstream_b200.workloads")

gb = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 2
wl = wls.make("cfg2", 77, gb=gb)
rq, rs = wl.directions
S = wl.nsessions
ctx = zk.Context(0)
ctx.set_timing(True)
device = "cuda:0"
d_req, d_rsp = dev.to_device(rq.data, device, pad=64), dev.to_device(rs.data, device, pad=64)
s_req = dev.DeviceSessions(rq.nbytes, rq.off, rq.flags, None, device)
s_rsp = dev.DeviceSessions(rs.nbytes, rs.off, rs.flags, None, device)
dxid = dev.DeviceXidTable(wl.layout, device)
c_req = dev.decode_dev_autosize(ctx, d_req, rq.nbytes, s_req, None, guess=rq.frames_hint, device=device)
dev.xidtab_build_dev(ctx, c_req, S, dxid)
c_rsp = dev.decode_dev_autosize(ctx, d_rsp, rs.nbytes, s_rsp, dxid, guess=rs.frames_hint, device=device)
cols = c_req.to_host()
rb, _ = zk.RequestBatch.from_columns(cols, rq.data)
dr = dev.DeviceRequests(rb, device)
out = None
res = {}
for _ in range(passes):
    dev.decode_dev(ctx, d_req, rq.nbytes, s_req, None, c_req)
    res["req"] = ctx.last_phase_ms()
    dev.xidtab_build_dev(ctx, c_req, S, dxid)
    dev.decode_dev(ctx, d_rsp, rs.nbytes, s_rsp, dxid, c_rsp)
    res["rsp"] = ctx.last_phase_ms()
    dev.session_stats_dev(ctx, c_rsp, S, device=device)
    o0, o1, o2, n, rc = dev.encode_dev(ctx, dr, rq.nbytes, device=device, out=out)
    out = (o0, o1, o2)
torch.cuda.synchronize()
print(json.dumps({"gb": gb, "sessions": S, "frames": [int(c_req.struct.n_frames), int(c_rsp.struct.n_frames)],
                  "phase_ms(split,emit,decode)": res, "encoded_bytes": n}))
ctx.close()
