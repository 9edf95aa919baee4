"""Development aid: where does the end-to-end (zkb_replay_host) time go?"""
import importlib, sys, time
import numpy as np, torch
sys.path.insert(0, ".")
zk = importlib.import_module("node-zkstream_b200")
synth = importlib.import_module("node-zkstream_b200.synth")
S = int(sys.argv[1]) if len(sys.argv) > 1 else 512
keep = []
def alloc(n):
    t = torch.empty(n + 64, dtype=torch.uint8, pin_memory=True); keep.append(t); return t.numpy()[:n]
pc = synth.native_cfg2(7, S, 20000, alloc=alloc)
ctx = zk.Context(0)
layout = zk.XidTable(np.zeros(S, np.int32), np.full(S, pc.frames, np.uint32))
tot = [0]
def light(first, ns, rq, rs): tot[0] += rq.n_frames + rs.n_frames
def heavy(first, ns, rq, rs):
    tot[0] += rq.n_frames + rs.n_frames + int((rs.meta & 0xff == 0).sum()) + rq.nbytes_columns()
nb = pc.req.size + pc.rsp.size
for name, cb, sl in (("light 256M", light, 0), ("heavy 256M", heavy, 0), ("light 128M", light, 128 << 20), ("light 512M", light, 512 << 20), ("light 1G", light, 1 << 30)):
    for rep in range(3):
        t0 = time.perf_counter()
        ctx.replay_host(pc.req, pc.req_off, pc.rsp, pc.rsp_off, layout.base, layout.count, cb,
                        req_flags=pc.req_flags, rsp_flags=pc.rsp_flags, copy=False, slice_bytes=sl)
        dt = time.perf_counter() - t0
    print("%-12s %.1f ms  %.1f GB/s wire" % (name, dt * 1e3, nb / dt / 1e9))
