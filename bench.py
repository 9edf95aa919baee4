#!/usr/bin/env python
"""bench.py — throughput of the batch ZooKeeper jute codec on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--config cfg1|cfg2|cfg3|cfg4|cfg5] [--gb G] [--damage P]

Workloads: the five shapes BASELINE.json names (node-zkstream_b200/workloads.py); the default,
and the one the headline metric is quoted on, is cfg2: a mixed request/response replay — per
connection a client->server stream and the paired server->client stream (CREATE /
GET_CHILDREN2 / SET_DATA / EXISTS / PING 20 % each, 5 % error replies, xids dense from 0, frame 0
= connect handshake), about 10 GB of wire bytes per GPU.  One STEP is one full pass of the hot
path over the batch:

    paired (cfg2, cfg5):  decode(request streams) -> build the xid->opcode table from them ->
                          decode(response streams, which resolve their opcodes through it)
    cfg1 / cfg3 / cfg4:   decode(the reply streams)

Before anything is timed, every rank checks what the GPU produced for ITS corpus against the CPU
oracle: all columns of a leading subsample row by row, and an order-sensitive 64-bit digest of every
column over the whole shard (oracle: zko_digest; GPU columns: verify.py).  A difference aborts the
run with a non-zero exit code.

`value` is whole-job decoded frames/s with the batch already resident in HBM (CUDA events on the
codec's stream, max over ranks).  `e2e` is the same pass through the host-buffer entry points of
the C ABI (zkb_replay_host for the paired workloads, zkb_decode_host otherwise: pinned host bytes
in, host columns out), all copies timed.  `roofline` describes the dominant kernel (k_decode).
`cpu_baseline` and `--impl reference` time the CPU port of the reference's decode loop
(oracle/zkoracle.c, faithful mode: 64 KiB chunks, per-frame copy + memmove as
lib/zk-streams.js:55-59) on the host cores; the reference itself is JavaScript and no node binary
exists on the box.  cfg5 adds the encode half: zkb_encode_dev / zkb_encode_host over every request
of the shard, with the oracle's encoder timed beside it.

Sessions are independent, so N GPUs decode N disjoint shards with no collective
("scaling": "weak": every rank gets the same amount of its own traffic).
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "decoded_zk_frames_per_sec"
UNIT = "frames/s"
# bytes per output row: this library's layout (include/zkbatch.h) and SURVEY.md §8d's accounting
ROW_BYTES = {"frames": 44, "stat": 72, "buf": 16, "acl": 8}
ROW_BYTES_SURVEY = {"frames": 36, "stat": 68, "buf": 12, "acl": 12, "vector_header": 12}


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            with open(p) as f:
                return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback 6650 (B200_PROFILING.md)"


def pcie_roofline(h2d_bytes, d2h_bytes, seconds, world):
    """e2e is bound by the host's PCIe complex.  Floor of a step = the larger of (upload bytes at the
    upload-only rate, download bytes at the download-only rate), both measured with `world` GPUs of
    the box copying at once (tools/pcie_peak.py -> profiles/pcie_peak_n<world>.json, else the single-GPU
    profiles/pcie_peak.json): the two directions are assumed to overlap perfectly.  achieved / peak are
    upload GB/s per GPU, measured and at that floor."""
    for name in ("pcie_peak_n%d.json" % world, "pcie_peak.json"):
        p = os.path.join(ROOT, "profiles", name)
        try:
            with open(p) as f:
                pk = json.load(f)
            up, down = float(pk["h2d_gbs"]), float(pk["d2h_gbs"])
        except Exception:
            continue
        floor = max(h2d_bytes / (up * 1e9), d2h_bytes / (down * 1e9))
        return {"bound": "pcie", "achieved": h2d_bytes / seconds / 1e9, "peak": h2d_bytes / floor / 1e9,
                "unit": "GB/s per GPU (upload)", "frac": floor / seconds,
                "limited_by": "upload" if h2d_bytes / up >= d2h_bytes / down else "download",
                "h2d_peak_gbs": up, "d2h_peak_gbs": down, "gpus_copying": pk.get("n_gpus", 1),
                "peak_source": "profiles/%s (measured on this pool)" % name}
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe).
    The sampler is started ahead of time (the process needs ~100 ms to come up) and polls every
    20 ms; stop(t0, t1) keeps the samples whose timestamps fall inside the timed window."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "20"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self, t0=None, t1=None):
        """t0, t1: datetime.now() taken right before / after the timed region (None = keep all)."""
        import datetime
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        rows = []
        try:
            with open(self.path) as f:
                for line in f:
                    p = [x.strip() for x in line.split(",")]
                    if len(p) < 9:
                        continue
                    try:
                        ts = datetime.datetime.strptime(p[0], "%Y/%m/%d %H:%M:%S.%f")
                        rows.append((ts, float(p[1]), float(p[2]), [v.lower().startswith("active") for v in p[5:9]]))
                    except ValueError:
                        continue
            os.unlink(self.path)
        except Exception:
            pass
        inside = [r for r in rows if t0 is not None and t0 <= r[0] <= t1]
        if inside:
            out["window"] = "timed region"
        else:
            # nothing landed inside a very short window: the samples around it are the same load
            inside = [r for r in rows if t0 is None or abs((r[0] - t0).total_seconds()) < 1.0 or
                      abs((r[0] - t1).total_seconds()) < 1.0]
            out["window"] = "within 1 s of the timed region"
        sm = sorted(r[1] for r in inside)
        reasons = set()
        for r in inside:
            out["sm_max_mhz"] = r[2]
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3]):
                if v:
                    reasons.add(name)
        if sm:
            out["sm_mhz"] = sm[len(sm) // 2]
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def gpu_index_for(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        ids = [x for x in vis.split(",") if x != ""]
        if local_rank < len(ids):
            return ids[local_rank]
    return str(local_rank)


def host_affinity():
    """Where this rank's host threads may run and which NUMA nodes the box shows (for the e2e numbers)."""
    out = {}
    try:
        cpus = sorted(os.sched_getaffinity(0))
        out["cpus"] = "%d-%d (%d)" % (cpus[0], cpus[-1], len(cpus)) if cpus else "none"
    except Exception:
        pass
    try:
        nodes = [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")]
        out["numa_nodes"] = len(nodes)
    except Exception:
        out["numa_nodes"] = None
    return out


# ---------------------------------------------------------------------------------------------
# CPU arms: the C port of the reference's decode loop (oracle/zkoracle.c) on the host cores


def _oracle():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import zkoracle
    return zkoracle


def cpu_pass(oracle, wl, mode, threads, nsess=None):
    """One pass of the CPU port over (the first `nsess` sessions of) every direction of `wl`.
    mode 1 = faithful (64 KiB chunks, per-frame copy + memmove), 0 = in-place walk.
    Returns (seconds, frames, bytes)."""
    zk = importlib.import_module("node-zkstream_b200")
    import numpy as np
    tot_t, tot_f, tot_b = 0.0, 0, 0
    for d in wl.directions:
        ns = d.nsessions if nsess is None else max(1, min(nsess, d.nsessions))
        off = d.off[: ns + 1]
        cut = int(off[-1])
        xt = None
        if d.xid == "paired":
            full = wl.truth_table()
            xt = zk.XidTable(full.base[:ns], full.count[:ns], full.ops[: int(full.count[:ns].sum())])
        elif d.xid is not None:
            xt = zk.XidTable(d.xid.base[:ns], d.xid.count[:ns], d.xid.ops[: int(d.xid.count[:ns].sum())])
        t, f = oracle.bench_decode(d.data[:cut], off, None if d.flags is None else d.flags[:ns],
                                   None if d.hs is None else d.hs[:ns], xt, mode, 65536, threads)
        tot_t += t
        tot_f += f
        tot_b += cut
    return tot_t, tot_f, tot_b


def cpu_decode_rate(oracle, wl, mode, threads, budget_s):
    """frames/s of the CPU port on a bounded sample of the workload (leading sessions of every
    direction), sized from a short calibration so that it runs for about `budget_s`."""
    ns0 = min(wl.nsessions, max(threads, 8))
    t, f, nb = cpu_pass(oracle, wl, mode, threads, ns0)
    ns = ns0
    want = int(ns0 * budget_s / max(t, 1e-3))
    if want > ns0 and wl.nsessions > ns0:
        ns = min(want, wl.nsessions)
        t, f, nb = cpu_pass(oracle, wl, mode, threads, ns)
    return f / t, ns, f, nb, t


def run_reference(args, rank, world):
    """--impl reference: the CPU port of the reference's decode loop on the host cores, on a bounded
    sample of the same workload."""
    if rank != 0:
        return
    wls = importlib.import_module("node-zkstream_b200.workloads")
    oracle = _oracle()
    threads = os.cpu_count() or 1
    wl = wls.sample(args.config, args.seed, threads)
    for _ in range(args.warmup):
        cpu_pass(oracle, wl, 1, threads)
    tot_t, tot_f = 0.0, 0
    for _ in range(args.steps):
        t, f, _ = cpu_pass(oracle, wl, 1, threads)
        tot_t += t
        tot_f += f
    value = tot_f / tot_t
    sample = "%d sessions (%.2f GB of %s) per step, faithful mode, 64 KiB chunks" % (
        wl.nsessions, wl.nbytes / 1e9, args.config)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": dict({"workload": wl.label + " (bounded sample)", "bytes_per_step": wl.nbytes}, **wl.info),
        "wire_GBps": wl.nbytes * args.steps / tot_t / 1e9,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference is JavaScript; no node binary on this box: oracle/zkoracle.c (C port of "
                "lib/zk-streams.js:39-100 + lib/zk-buffer.js readers) timed instead",
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# parity gate: what the GPU produced for this rank's corpus against the CPU oracle


def parity_gate(oracle, zk, dev, ver, wl, dcols, dxid, threads, sub_bytes=192 << 20):
    """Returns the `parity` object of the bench line; ok = False on any difference.
    (1) subsample: the leading sessions of every direction (or, for one giant session, a prefix cut at
        a frame boundary) decoded by the oracle, every column compared row by row with the GPU's rows;
    (2) whole shard: oracle digest of every column (zko_digest, all host threads) against the digest
        of the GPU's device-resident columns."""
    import numpy as np
    import torch
    out = {"ok": True, "sessions_checked": 0, "frames_checked": 0, "columns_compared": 0, "digest_ok": True,
           "digest_columns": 0, "frames_digested": 0, "problems": []}
    t0 = time.time()
    want_req, ns_req = None, 0
    for di, (d, dc) in enumerate(zip(wl.directions, dcols)):
        # ---- (1) subsample ----
        ns = int(np.searchsorted(d.off, sub_bytes, "right")) - 1
        whole = ns >= 1
        if whole:
            ns = min(ns, 64, d.nsessions)
            if d.xid == "paired":
                ns = min(ns, ns_req)         # the oracle's table comes from the same connections' requests
            cut = int(d.off[ns])
            rows = None
        else:
            # one session larger than the sample: cut it at a frame boundary the GPU found
            fo = dc.t["frame_off"][: int(dc.struct.n_frames)]
            rows = int(torch.searchsorted(fo, torch.tensor([sub_bytes], device=fo.device, dtype=fo.dtype)).item())
            rows = max(1, min(rows, int(dc.struct.n_frames) - 1))
            cut = int(fo[rows].item())
            ns = 1
        xt = None
        if d.xid == "paired":
            lay = zk.XidTable(wl.layout.base[:ns], wl.layout.count[:ns])
            xt = oracle.xidtab_build(want_req, lay)
        elif d.xid is not None:
            xt = zk.XidTable(d.xid.base[:ns], d.xid.count[:ns], d.xid.ops[: int(d.xid.count[:ns].sum())])
        off = d.off[: ns + 1].copy()
        off[-1] = cut
        want = oracle.decode(d.data[:cut], off, None if d.flags is None else d.flags[:ns],
                             None if d.hs is None else d.hs[:ns], xt)
        if di == 0:
            want_req, ns_req = want, ns
        for cnt, cols in (("n_frames", zk._abi.FRAME_COLS), ("n_stat", zk._abi.STAT_COLS),
                          ("n_buf", zk._abi.BUF_COLS), ("n_acl", zk._abi.ACL_COLS)):
            n = int(getattr(want, cnt))
            for name, dt in cols:
                got = dc.t[name][:n].cpu().numpy().view(dev._NP_DT[dt])
                out["columns_compared"] += 1
                if not np.array_equal(got, getattr(want, name)):
                    bad = np.flatnonzero(got != getattr(want, name))
                    out["problems"].append("%s: column %s differs from the oracle, first at row %d" % (d.name, name, bad[0]))
        if whole:
            for name, dt in zk._abi.SESS_COLS:
                n = ns if name != "sess_first_frame" else ns + 1
                got = dc.t[name][:n].cpu().numpy().view(dev._NP_DT[dt])
                out["columns_compared"] += 1
                if not np.array_equal(got, getattr(want, name)[:n]):
                    out["problems"].append("%s: column %s differs from the oracle" % (d.name, name))
        out["sessions_checked"] += ns
        out["frames_checked"] += int(want.n_frames)
        # ---- (2) digest of the whole shard ----
        if d.xid == "paired":
            xt_full = wl.truth_table()
            built = dxid.ops[: xt_full.nops].cpu().numpy().view(np.int8)
            if not np.array_equal(built, xt_full.ops):
                out["problems"].append("xid->opcode table built on the GPU differs from the generator's")
        else:
            xt_full = d.xid
        odig, ocnt = oracle.digest(d.data, d.off, d.flags, d.hs, xt_full, threads)
        gcnt = tuple(int(getattr(dc.struct, k)) for k in ("n_frames", "n_stat", "n_buf", "n_acl"))
        if gcnt != ocnt:
            out["problems"].append("%s: row counts %r, oracle %r" % (d.name, gcnt, ocnt))
            out["digest_ok"] = False
        else:
            gdig = ver.digest_device_columns(dc)
            bad = [k for k in ver.COLUMNS if gdig[k] != odig[k]]
            out["digest_columns"] += len(ver.COLUMNS)
            out["frames_digested"] += ocnt[0]
            if bad:
                out["digest_ok"] = False
                out["problems"].append("%s: digest differs for %s" % (d.name, ", ".join(bad)))
    out["ok"] = not out["problems"]
    out["seconds"] = round(time.time() - t0, 2)
    out["how"] = ("all columns row by row on the leading sessions (or a frame-aligned prefix of a giant session); "
                  "mix64(value)*(row+1) digest of every column over the whole shard, oracle vs GPU columns")
    if not out["problems"]:
        del out["problems"]
    return out


# ---------------------------------------------------------------------------------------------
# the GPU arm


def damage_replies(wl, dcols, d_bufs, rate, torch, np):
    """--damage P: make one GET_CHILDREN2 reply per 1/P frames lie about its child count (count + 1:
    the walk runs into the Stat, the frame becomes BAD_DECODE and everything the splitter predicted
    for it is wrong), in the host copy and in the device copy.  Returns the number of frames damaged."""
    d = wl.directions[-1]
    dc = dcols[-1]
    nf = int(dc.struct.n_frames)
    meta = dc.t["meta"][:nf]
    err = dc.t["err"][:nf]
    cand = torch.nonzero(((meta & 0xffff) == (12 << 8)) & (err == 0) & (((meta >> 16) & 0xff) == 0)).flatten()
    want = max(1, int(nf * rate))
    if cand.numel() == 0:
        return 0
    pick = cand[torch.linspace(0, cand.numel() - 1, min(want, cand.numel())).long()]
    pos = (dc.t["frame_off"][pick] + 4 + 16 + 3).cpu().numpy()          # low byte of the count
    d.data[pos] = d.data[pos] + 1
    d_bufs[-1][torch.from_numpy(pos).to(d_bufs[-1].device)] += 1
    return int(pick.numel())


def run_ours(args, rank, local_rank, world):
    import datetime
    import numpy as np
    import torch
    zk = importlib.import_module("node-zkstream_b200")
    dev = importlib.import_module("node-zkstream_b200.device")
    ver = importlib.import_module("node-zkstream_b200.verify")
    wls = importlib.import_module("node-zkstream_b200.workloads")

    use_dist = world > 1
    if use_dist:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    device = "cuda:%d" % local_rank
    ctx = zk.Context(local_rank)
    ctx.set_timing(True)
    ext = torch.cuda.ExternalStream(ctx.stream, device=device)
    threads = max(1, (os.cpu_count() or 1) // max(world, 1)) if use_dist else (os.cpu_count() or 1)

    # ---- workload: bounded by host memory ----
    gb = args.gb if args.gb is not None else wls.DEFAULT_GB[args.config]
    try:
        import psutil
        avail = psutil.virtual_memory().available / max(world, 1)
        gb = min(gb, avail * 0.40 / 1e9)
    except Exception:
        pass
    t0 = time.time()
    wl = wls.make(args.config, args.seed + 1000003 * rank, gb=gb, pinned=True)
    gen_s = time.time() - t0
    dirs = wl.directions
    S = wl.nsessions

    # ---- device-resident copies of the batch (outside every timed region) ----
    d_bufs = [dev.to_device(d.data, device, pad=64) for d in dirs]
    d_sess = [dev.DeviceSessions(d.nbytes, d.off, d.flags, d.hs, device) for d in dirs]
    dxid = None
    d_xids = []
    for d in dirs:
        if d.xid == "paired":
            dxid = dev.DeviceXidTable(wl.layout, device)
            d_xids.append(dxid)
        elif d.xid is not None:
            d_xids.append(dev.DeviceXidTable(d.xid, device))
        else:
            d_xids.append(None)
    dcols = []
    for i, d in enumerate(dirs):
        dcols.append(dev.decode_dev_autosize(ctx, d_bufs[i], d.nbytes, d_sess[i], d_xids[i], guess=d.frames_hint,
                                             device=device))
        if wl.paired and i == 0:
            dev.xidtab_build_dev(ctx, dcols[0], S, dxid)
    torch.cuda.synchronize()

    slow = {}

    def step_dev(acc=None):
        for i, d in enumerate(dirs):
            rc = dev.decode_dev(ctx, d_bufs[i], d.nbytes, d_sess[i], d_xids[i], dcols[i])
            assert rc == 0, "zkb_decode_dev rc %d" % rc
            if acc is not None:
                acc[i].append(ctx.last_phase_ms())
            slow[i] = ctx.last_slow_frames
            if wl.paired and i == 0:
                dev.xidtab_build_dev(ctx, dcols[0], S, dxid)

    def barrier():
        if use_dist:
            dist.barrier()
        torch.cuda.synchronize()

    damaged = 0
    if args.damage > 0:
        damaged = damage_replies(wl, dcols, d_bufs, args.damage, torch, np)
        dcols[-1] = dev.decode_dev_autosize(ctx, d_bufs[-1], dirs[-1].nbytes, d_sess[-1], d_xids[-1],
                                            guess=dirs[-1].frames_hint, device=device)
    mis0 = ctx.row_mispredicts
    for _ in range(args.warmup):
        step_dev()

    # ---- parity gate on exactly this corpus, before anything is timed ----
    parity = None
    if not args.no_parity:
        parity = parity_gate(_oracle(), zk, dev, ver, wl, dcols, dxid, threads)
        ok = torch.tensor([1 if parity["ok"] else 0], device=device)
        if use_dist:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if not parity["ok"]:
            print("rank %d: PARITY FAILED: %s" % (rank, "; ".join(parity.get("problems", []))), file=sys.stderr, flush=True)
        if int(ok.item()) == 0:
            if use_dist:
                dist.destroy_process_group()
            sys.exit(3)

    counts = []
    for i, dc in enumerate(dcols):
        s = dc.struct
        nf = int(s.n_frames)
        meta = dc.t["meta"][:nf]
        okf = (meta & 0xff) == 0
        op = (meta >> 8) & 0xff
        kind = (meta >> 16) & 0xff
        if dirs[i].flags is not None and int(dirs[i].flags[0]) & 1:
            vec = int((okf & (kind == 1) & (op == 1)).sum().item()) + 3 * int((okf & (kind == 1) & (op == 101)).sum().item())
        else:
            e0 = dc.t["err"][:nf] == 0
            vec = int((okf & (kind == 0) & e0 & ((op == 8) | (op == 12) | (op == 6))).sum().item())
        counts.append(dict(frames=nf, stat=int(s.n_stat), buf=int(s.n_buf), acl=int(s.n_acl), vectors=vec,
                           slow_frames=slow.get(i), bad_decode=int((~okf).sum().item())))
    frames_per_step = sum(c["frames"] for c in counts)

    sampler = ClockSampler(gpu_index_for(local_rank))
    sampler.start()
    for _ in range(2):          # keeps the GPU under the same load while the sampler comes up
        step_dev()
    phases = [[] for _ in dirs]
    launches0 = ctx.kernel_launches
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    wall0 = datetime.datetime.now()
    e0.record(ext)
    for _ in range(args.steps):
        step_dev(phases)
    e1.record(ext)
    barrier()
    wall1 = datetime.datetime.now()
    clocks = sampler.stop(wall0, wall1)
    launches = ctx.kernel_launches - launches0
    ms = e0.elapsed_time(e1)
    if use_dist:
        t = torch.tensor([ms], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        fr = torch.tensor([frames_per_step, wl.nbytes], device=device, dtype=torch.float64)
        dist.all_reduce(fr, op=dist.ReduceOp.SUM)
        total_frames, total_bytes = float(fr[0].item()), float(fr[1].item())
    else:
        total_frames, total_bytes = float(frames_per_step), float(wl.nbytes)
    ms_per_step = ms / args.steps
    value = total_frames / (ms_per_step * 1e-3)

    # ---- roofline of the dominant kernel (k_decode), from the library's own CUDA events ----
    def out_bytes(c, rb):
        return (c["frames"] * rb["frames"] + c["stat"] * rb["stat"] + c["buf"] * rb["buf"] + c["acl"] * rb["acl"] +
                c["vectors"] * rb.get("vector_header", 0))

    alg = [d.nbytes + out_bytes(c, ROW_BYTES) for d, c in zip(dirs, counts)]
    alg_survey = [d.nbytes + out_bytes(c, ROW_BYTES_SURVEY) for d, c in zip(dirs, counts)]
    dec_ms = [sum(p[2] for p in ph) / max(1, len(ph)) for ph in phases]
    split_ms = sum(sum(p[0] + p[1] for p in ph) for ph in phases) / max(1, args.steps)
    peak, peak_src = measured_peak()
    ach = sum(alg) / (sum(dec_ms) * 1e-3) / 1e9
    ach_survey = sum(alg_survey) / (sum(dec_ms) * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "k_decode_traffic.json")
    if os.path.exists(tp):
        try:
            with open(tp) as f:
                tj = json.load(f)
            if tj.get("config", "cfg2") == args.config and tj.get("sessions") == S:
                traffic = tj.get("dram_bytes_per_launch")
        except Exception:
            pass
    roofline = {
        "bound": "hbm", "kernel": "k_decode", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
        "peak_source": peak_src,
        "frac_of_nominal_8TBps": ach / 8000.0,
        "frac_survey_bytes": ach_survey / peak,
        "traffic": traffic,
        "algorithmic_bytes_per_launch": sum(alg) / len(alg),
        "algorithmic_bytes_per_frame": sum(alg) / max(1, frames_per_step),
        "algorithmic_bytes_per_frame_survey": sum(alg_survey) / max(1, frames_per_step),
        "launch_ms": {d.name: m for d, m in zip(dirs, dec_ms)},
        "share_of_step": sum(dec_ms) / ms_per_step if not use_dist else None,
        "split_ms_per_step": split_ms,
        "step_frac": sum(alg) / (ms_per_step * 1e-3) / 1e9 / peak if not use_dist else None,
        "accounting": "wire bytes read once + rows written once (44 B FRAMES, 72 STAT, 16 BUF, 8 ACL); "
                      "frac_survey_bytes uses SURVEY 8d's 36/68/12 per descriptor, 12 per vector header",
    }

    # ---- end to end through the host-buffer entry points ----
    e2e_steps = max(2, min(args.steps, 4))
    h2d = d2h = 0
    dev_sums = None
    if wl.paired:
        # what the e2e pass must deliver, from the device-resident columns already checked above
        dev_sums = [int(dcols[1].t["zxid"][: counts[1]["frames"]].sum().item()),
                    int((dcols[1].t["meta"][: counts[1]["frames"]].to(torch.int64) & 0xFFFFFFFF).sum().item())]

    if wl.paired and not args.e2e_full_columns:
        # the replay's compact results (derivable columns left out of the download): what a replay
        # tool on a multi-GPU host would use, because the download is what such a host runs out of first
        ctx.set_replay_compact(True)

    def step_e2e(check=False):
        nonlocal h2d, d2h
        tot = {"frames": 0, "d2h": 0, "ok": 0, "zxid": 0, "meta": 0}
        if wl.paired:
            rq, rs = dirs

            def on_slice(first, ns, a, b):
                # the "result read": every slice's columns are on the host; read one status word per page
                tot["frames"] += a.n_frames + b.n_frames
                tot["d2h"] += a.nbytes_columns() + b.nbytes_columns()
                tot["ok"] += int(b.meta[::1024].sum(dtype=np.int64))
                if check:
                    tot["zxid"] += int(b.zxid.sum(dtype=np.int64))
                    tot["meta"] += int(b.meta.sum(dtype=np.int64))

            ctx.replay_host(rq.data, rq.off, rs.data, rs.off, wl.layout.base, wl.layout.count, on_slice,
                            req_flags=rq.flags, rsp_flags=rs.flags, copy=False)
            h2d = wl.nbytes + 2 * (S + 1) * 8 + 2 * S + 16 * S
        else:
            d = dirs[0]
            cols = ctx.decode_host(d.data, d.off, d.flags, d.hs, d.xid, copy=False)
            tot["frames"] = cols.n_frames
            tot["d2h"] = cols.nbytes_columns()
            tot["ok"] = int(cols.meta[::1024].sum(dtype=np.int64))       # one status word per page
            h2d = d.nbytes + (S + 1) * 8 + (d.xid.nops + 16 * S if d.xid is not None else 0)
        d2h = tot["d2h"]
        if check and dev_sums is not None:
            M = (1 << 64) - 1
            assert (tot["zxid"] & M) == (dev_sums[0] & M) and tot["meta"] == dev_sums[1], \
                "e2e pass delivered different rows than the device-resident pass"
        return tot["frames"]

    got = step_e2e(check=True)
    assert got == frames_per_step, "e2e pass decoded %d frames, device-resident pass %d" % (got, frames_per_step)
    barrier()
    e0.record(ext)
    t_wall = time.perf_counter()
    for _ in range(e2e_steps):
        got = step_e2e()
    e1.record(ext)
    barrier()
    t_wall = time.perf_counter() - t_wall
    e2e_ms = max(e0.elapsed_time(e1), t_wall * 1e3)
    if use_dist:
        t = torch.tensor([e2e_ms], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    assert got == frames_per_step
    e2e_value = total_frames / (e2e_ms / e2e_steps * 1e-3)
    if parity is not None:
        parity["e2e_rows_checked"] = bool(wl.paired)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": dict({"workload": wl.label, "config": args.config,
                        "sessions_per_gpu": S, "wire_bytes_per_gpu": wl.nbytes, "frames_per_gpu": frames_per_step,
                        "l2": "inputs (%.2f GB per GPU) larger than the 126 MB L2; no flush needed" % (wl.nbytes / 1e9),
                        "parallelism": "sessions sharded over %d GPU(s), no collective" % world,
                        "seed": args.seed, "corpus_gen_s": round(gen_s, 1)}, **wl.info),
        "wire_GBps": total_bytes / (ms_per_step * 1e-3) / 1e9,
        "rows_per_gpu": {d.name: c for d, c in zip(dirs, counts)},
        "parity": parity,
        "roofline": roofline,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps,
                "api": ("zkb_replay_host" + ("" if args.e2e_full_columns else " (compact results: zkb_ctx_set_replay_compact)"))
                if wl.paired else "zkb_decode_host",
                "wire_GBps": total_bytes / (e2e_ms / e2e_steps * 1e-3) / 1e9,
                "pcie_roofline": pcie_roofline(h2d, d2h, e2e_ms / e2e_steps * 1e-3, world),
                "host": host_affinity()},
        "gpu_launches": int(launches),
        "split_fixups": ctx.split_fixups,
        "row_mispredicts": ctx.row_mispredicts - mis0,
        "clocks": clocks,
    }
    if args.damage > 0:
        line["damage"] = {"rate": args.damage, "frames_damaged": damaged,
                          "how": "GET_CHILDREN2 replies whose child count was raised by one"}

    # ---- the encode half of the path ----
    if not args.no_encode and wl.paired and (args.config == "cfg5" or (rank == 0 and world == 1)):
        line["encode"] = encode_leg(args, zk, dev, ctx, wl, dcols[0], device, ext, torch, np,
                                    full=(args.config == "cfg5"), cpu=(rank == 0 and not args.no_cpu))

    # ---- CPU baseline beside it (rank 0, N=1 only) ----
    if rank == 0 and world == 1 and not args.no_cpu:
        oracle = _oracle()
        cores = os.cpu_count() or 1
        # the CPU port shards sessions over threads: a one-session workload is timed on the same
        # frames cut into enough sessions for every core
        cwl = wl if wl.nsessions >= 2 * cores else wls.sample(args.config, args.seed, cores)
        rate, ns, f, nb, t = cpu_decode_rate(oracle, cwl, 1, cores, 8.0)
        rate_ip, ns2, f2, nb2, t2 = cpu_decode_rate(oracle, cwl, 0, cores, 4.0)
        rate_1, ns3, f3, nb3, t3 = cpu_decode_rate(oracle, cwl, 1, 1, 3.0)
        S = cwl.nsessions
        line["cpu_baseline"] = {
            "value": rate, "unit": UNIT, "cores": min(cores, ns), "kind": "port",
            "sample": "first %d of %d sessions, every direction (%.2f GB, %d frames, %.1f s): C port of the "
                      "reference loop, faithful mode (64 KiB chunks, per-frame copy + memmove)" % (ns, S, nb / 1e9, f, t),
            "in_place_value": rate_ip, "in_place_sample": "%d sessions, %.1f s, no per-frame copies" % (ns2, t2),
            "single_thread_value": rate_1,
            "e2e_over_faithful": e2e_value / rate, "e2e_over_in_place": e2e_value / rate_ip,
            "node": "unavailable on this box (no node binary, no network)",
        }
    if rank == 0:
        print(json.dumps(line), flush=True)
    if use_dist:
        dist.destroy_process_group()
    ctx.close()


def encode_leg(args, zk, dev, ctx, wl, dc_req, device, ext, torch, np, full, cpu):
    """The encode half (lib/zk-streams.js:121-148, lib/zk-buffer.js:97-136): the decoded requests of
    this rank's shard (cfg5: all of them; otherwise the first 160 connections) turned back into
    descriptors and re-encoded — device-resident (zkb_encode_dev), end to end from host descriptors
    to host frames (zkb_encode_host), and by the oracle's encoder on the host cores."""
    req = wl.directions[0]
    ns = req.nsessions if full else min(req.nsessions, 160)
    cut = int(req.off[ns])
    if full:
        cols = dc_req.to_host()
    else:
        cols = ctx.decode_host(req.data[:cut], req.off[: ns + 1], req.flags[:ns], None, None)
    rb, _ = zk.RequestBatch.from_columns(cols, req.data[:cut])
    del cols
    dr = dev.DeviceRequests(rb, device)
    out = None
    for _ in range(3):
        o0, o1, o2, enc_n, rc = dev.encode_dev(ctx, dr, cut, device=device, out=out)
        out = (o0, o1, o2)
        assert rc == 0
    l0 = ctx.kernel_launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    K = 5
    with torch.cuda.stream(ext):
        ev0.record()
        for _ in range(K):
            dev.encode_dev(ctx, dr, cut, device=device, out=out)
        ev1.record()
    torch.cuda.synchronize()
    enc_ms = ev0.elapsed_time(ev1) / K
    launches = (ctx.kernel_launches - l0) / K
    payload = sum(int(np.maximum(rb.cols[k], 0).sum()) for k in ("path_len", "data_len", "elem_len",
                                                                "acl_scheme_len", "acl_id_len"))
    alg = dr.nbytes() - int(rb.arena.size) - 32 + payload + enc_n + 9 * rb.n
    peak = measured_peak()[0]
    res = {
        "packets": rb.n, "frame_bytes": enc_n, "ms": enc_ms, "packets_per_s": rb.n / (enc_ms * 1e-3),
        "frames_GBps": enc_n / (enc_ms * 1e-3) / 1e9,
        "roofline": {"bound": "hbm", "kernel": "k_enc_write (+ size/scan/span kernels)",
                     "achieved": alg / (enc_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": alg / (enc_ms * 1e-3) / 1e9 / peak,
                     "algorithmic_bytes": "descriptor columns + payload bytes read + frames, frame_off and status written"},
        "launches_per_call": launches,
        "sample": ("every request of the shard" if full else "requests of the first %d connections" % ns) +
                  ", re-encoded from their decoded columns",
    }
    # parity: the GPU's frames against the oracle's on a bounded leading part of the same batch
    if cpu:
        oracle = _oracle()
        nchk = min(rb.n, 2_000_000)
        sub = zk.RequestBatch.__new__(zk.RequestBatch)
        sub.n, sub.arena, sub.names = nchk, rb.arena, None
        sub.cols = dict(rb.cols)
        for name, _ in zk._abi.REQ_COLS:
            sub.cols[name] = rb.cols[name][:nchk]
        sub.n_elem, sub.n_acl = rb.n_elem, rb.n_acl
        od, oo, os_ = oracle.encode(sub)
        g_off = out[1][: nchk + 1].cpu().numpy().view(np.uint64)
        g_bytes = out[0][: int(g_off[-1])].cpu().numpy()
        res["parity"] = {"packets_checked": nchk,
                         "ok": bool(np.array_equal(g_off, oo) and np.array_equal(g_bytes, od) and
                                    np.array_equal(out[2][:nchk].cpu().numpy(), os_))}
        assert res["parity"]["ok"], "encoded frames differ from the oracle's"
        threads = os.cpu_count() or 1
        t, nb = oracle.bench_encode(sub, threads)
        t1, _ = oracle.bench_encode(sub, 1)
        res["cpu_baseline"] = {"value": nchk / t, "unit": "packets/s", "cores": threads, "kind": "port",
                               "single_thread_value": nchk / t1,
                               "sample": "first %d packets of the same batch, oracle/zkoracle.c encoder (%.2f s)" % (nchk, t)}
    # end to end: host descriptors + arena in, host frames out
    if full or cpu:
        import ctypes as C

        def host_call():
            r, keep = rb.struct()
            o, fo, st, n = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_uint64()
            rc = ctx._lib.zkb_encode_host(ctx._h, C.byref(r), C.byref(o), C.byref(fo), C.byref(st), C.byref(n))
            assert rc == 0 and n.value == enc_n
            del keep

        host_call()                      # first call sizes the context's staging buffers
        reps = 2
        t0 = time.perf_counter()
        for _ in range(reps):
            host_call()
        dt = (time.perf_counter() - t0) / reps
        res["e2e"] = {"value": rb.n / dt, "unit": "packets/s", "ms": dt * 1e3, "api": "zkb_encode_host",
                      "h2d_bytes": int(dr.nbytes()), "d2h_bytes": int(enc_n + 9 * rb.n),
                      "note": "pageable numpy descriptors in, context-owned pinned frames out"}
    del dr, out
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg2", choices=["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"])
    ap.add_argument("--gb", type=float, default=None, help="wire bytes per GPU, in GB (default: the size BASELINE names)")
    ap.add_argument("--damage", type=float, default=0.0, help="fraction of frames given a lying child count (paired configs)")
    ap.add_argument("--seed", type=int, default=20240607)
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline legs")
    ap.add_argument("--no-encode", action="store_true", help="skip the encode leg")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle parity gate (development only)")
    ap.add_argument("--e2e-full-columns", action="store_true",
                    help="e2e replay downloads every column instead of the compact results")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = env_int("RANK", 0)
    local_rank = env_int("LOCAL_RANK", 0)
    world = env_int("WORLD_SIZE", 1)
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 1000)] + sys.argv
        os.execv(sys.executable, cmd)
    run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
