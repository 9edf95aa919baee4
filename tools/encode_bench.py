"""Development aid: device-resident encode throughput on the request mix of the bench workload.

    python tools/encode_bench.py [sessions] [frames_per_session]

Decodes the request streams, turns the columns back into descriptors (RequestBatch.from_columns:
the arena is the original capture), uploads them and times zkb_encode_dev with CUDA events."""
import importlib
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
zk = importlib.import_module("node-zkstream_b200")
synth = importlib.import_module("node-zkstream_b200.synth")
dev = importlib.import_module("node-zkstream_b200.device")

S = int(sys.argv[1]) if len(sys.argv) > 1 else 320
F = int(sys.argv[2]) if len(sys.argv) > 2 else 20000
pc = synth.native_cfg2(41, S, F)
ctx = zk.Context(0)
cols = ctx.decode_host(pc.req, pc.req_off, pc.req_flags, None, None)
rb, idx = zk.RequestBatch.from_columns(cols, pc.req)
dr = dev.DeviceRequests(rb)
cap = int(pc.req.size)
out = None
for _ in range(3):
    o0, o1, o2, n, rc = dev.encode_dev(ctx, dr, cap, out=out)
    out = (o0, o1, o2)
assert rc == 0
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
torch.cuda.synchronize()
K = 10
l0 = ctx.kernel_launches
ev[0].record()
for _ in range(K):
    dev.encode_dev(ctx, dr, cap, out=out)
ev[1].record()
torch.cuda.synchronize()
ms = ev[0].elapsed_time(ev[1]) / K
desc = sum(t.numel() for k, t in dr.t.items() if k != "arena")
payload = int(np.maximum(rb.cols["path_len"], 0).sum() + np.maximum(rb.cols["data_len"], 0).sum() +
              np.maximum(rb.cols["elem_len"], 0).sum() + np.maximum(rb.cols["acl_scheme_len"], 0).sum() +
              np.maximum(rb.cols["acl_id_len"], 0).sum())
alg = desc + payload + n + 8 * (rb.n + 1) + rb.n
print(json.dumps({"packets": rb.n, "wire_bytes": n, "ms": ms, "packets_per_s": rb.n / ms * 1e3,
                  "wire_GBps": n / ms / 1e6, "algorithmic_GBps": alg / ms / 1e6,
                  "frac_of_hbm_peak_6539.5": alg / ms / 1e6 / 6539.5, "launches_per_call": (ctx.kernel_launches - l0) / K,
                  "note": "torch events on the default stream bracket K synchronous calls; each call ends with a stream sync"}))
