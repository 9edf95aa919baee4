#!/usr/bin/env python
"""bench.py — throughput of the batch ZooKeeper jute decoder on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[1], SURVEY.md §8d "cfg2"): a mixed request/response
replay — per session a client->server stream and the paired server->client stream
(CREATE / GET_CHILDREN2 / SET_DATA / EXISTS / PING 20 % each, 5 % error replies, xids
dense from 0, frame 0 = connect handshake), about 10 GB of wire bytes per GPU.  One STEP
is one full pass of the hot path over that batch:

    decode(request streams) -> build the xid->opcode table from them -> decode(response
    streams, which resolve their opcodes through that table)

`value` is whole-job decoded frames/s with the batch already resident in HBM (CUDA events
on the codec's stream, max over ranks).  `e2e` is the same pass through the host-buffer
entry point of the C ABI (zkb_replay_host: pinned host bytes in, host columns out, slices
pipelined over upload / decode / download streams), all copies timed.
`roofline` describes the dominant kernel (k_decode).  `cpu_baseline` and
`--impl reference` time the CPU port of the reference's decode loop (oracle/zkoracle.c,
faithful mode: 64 KiB chunks, per-frame copy + memmove as lib/zk-streams.js:55-59) on the
host cores; the reference itself is JavaScript and no node binary exists on the box.

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
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "decoded_zk_frames_per_sec"
UNIT = "frames/s"
FRAMES_PER_SESSION = 20000
BYTES_PER_SESSION = 7.79e6          # request + response stream of one session (measured)
# bytes k_decode writes per row: FRAMES (all nine columns, 44 B), STAT, BUF, ACL
ROW_BYTES = {"frames_decode": 44, "frames_emit": 0, "stat": 72, "buf": 16, "acl": 8}


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
                return float(json.load(f)["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


def pcie_roofline(h2d_gbps_per_gpu):
    """e2e is bound by the upload: fraction of the measured duplex PCIe rate (profiles/pcie_peak.json)."""
    p = os.path.join(ROOT, "profiles", "pcie_peak.json")
    try:
        with open(p) as f:
            peak = float(json.load(f)["duplex_h2d_equiv_gbs"])
        return {"bound": "pcie", "achieved": h2d_gbps_per_gpu, "peak": peak, "unit": "GB/s per GPU (upload)",
                "frac": h2d_gbps_per_gpu / peak, "peak_source": "profiles/pcie_peak.json (measured on this pool)"}
    except Exception:
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


def make_corpus(synth, seed, nsessions, pinned):
    import numpy as np
    if pinned:
        import torch
        keep = []

        def alloc(n):
            t = torch.empty(n + 64, dtype=torch.uint8, pin_memory=True)
            keep.append(t)
            return t.numpy()[:n]
    else:
        keep = None
        alloc = None
    pc = synth.native_cfg2(seed, nsessions, FRAMES_PER_SESSION, alloc=alloc)
    pc._keep = keep
    return pc


def cpu_decode_rate(oracle, zk, pc, nsessions, mode, threads, budget_s):
    """frames/s of the CPU port on a bounded sample of the workload (first sessions of both
    directions), sized from a short calibration so that it runs for about `budget_s`."""
    import numpy as np
    xt_full = pc.xidtable()

    def run(ns):
        ns = max(1, min(ns, nsessions))
        roff = pc.req_off[: ns + 1]
        soff = pc.rsp_off[: ns + 1]
        xt = zk.XidTable(np.zeros(ns, np.int32), np.full(ns, pc.frames, np.uint32), xt_full.ops[: ns * pc.frames])
        t1, f1 = oracle.bench_decode(pc.req[: int(roff[-1])], roff, pc.req_flags[:ns], None, None, mode, 65536, threads)
        t2, f2 = oracle.bench_decode(pc.rsp[: int(soff[-1])], soff, pc.rsp_flags[:ns], None, xt, mode, 65536, threads)
        return t1 + t2, f1 + f2, ns, int(roff[-1]) + int(soff[-1])

    t, f, ns, nb = run(max(threads, 8))
    rate = f / t
    want = int(ns * budget_s / max(t, 1e-3))
    if want > ns:
        t, f, ns, nb = run(want)
        rate = f / t
    return rate, ns, f, nb, t


def run_reference(args, rank, world):
    """--impl reference: the CPU port of the reference's decode loop on the host cores."""
    if rank != 0:
        return
    zk = importlib.import_module("node-zkstream_b200")
    synth = importlib.import_module("node-zkstream_b200.synth")
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import zkoracle as oracle
    import numpy as np
    threads = os.cpu_count() or 1
    nsessions = max(threads * 4, 64)
    pc = make_corpus(synth, args.seed, nsessions, pinned=False)
    xt = pc.xidtable()

    def step():
        t1, f1 = oracle.bench_decode(pc.req, pc.req_off, pc.req_flags, None, None, 1, 65536, threads)
        t2, f2 = oracle.bench_decode(pc.rsp, pc.rsp_off, pc.rsp_flags, None, xt, 1, 65536, threads)
        return t1 + t2, f1 + f2

    for _ in range(args.warmup):
        step()
    tot_t, tot_f = 0.0, 0
    for _ in range(args.steps):
        t, f = step()
        tot_t += t
        tot_f += f
    value = tot_f / tot_t
    nbytes = int(pc.req.size + pc.rsp.size)
    sample = "%d sessions x %d frames (%.2f GB of the cfg2 replay) per step, faithful mode, 64 KiB chunks" % (
        nsessions, FRAMES_PER_SESSION, nbytes / 1e9)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "cfg2 mixed request/response replay (bounded sample)", "sessions": nsessions,
                   "frames_per_session": FRAMES_PER_SESSION, "bytes_per_step": nbytes},
        "wire_GBps": nbytes * args.steps / tot_t / 1e9,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference is JavaScript; no node binary on this box: oracle/zkoracle.c (C port of "
                "lib/zk-streams.js:39-100 + lib/zk-buffer.js readers) timed instead",
    }
    print(json.dumps(line), flush=True)


def run_ours(args, rank, local_rank, world):
    import numpy as np
    import torch
    zk = importlib.import_module("node-zkstream_b200")
    synth = importlib.import_module("node-zkstream_b200.synth")
    dev = importlib.import_module("node-zkstream_b200.device")

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

    # ---- workload: ~args.gb GB of wire bytes per GPU, bounded by host memory ----
    nsessions = max(16, int(args.gb * 1e9 / BYTES_PER_SESSION))
    try:
        import psutil
        avail = psutil.virtual_memory().available / max(world, 1)
        fit = int(avail * 0.45 / BYTES_PER_SESSION)
        if fit < nsessions:
            nsessions = max(16, fit)
    except Exception:
        pass
    t0 = time.time()
    pc = make_corpus(synth, args.seed + 1000003 * rank, nsessions, pinned=True)
    gen_s = time.time() - t0
    nreq, nrsp = int(pc.req.size), int(pc.rsp.size)
    S = nsessions

    # ---- device-resident copies of the batch (outside every timed region) ----
    d_req = dev.to_device(pc.req, device, pad=64)
    d_rsp = dev.to_device(pc.rsp, device, pad=64)
    ds_req = dev.DeviceSessions(nreq, pc.req_off, pc.req_flags, None, device)
    ds_rsp = dev.DeviceSessions(nrsp, pc.rsp_off, pc.rsp_flags, None, device)
    layout = zk.XidTable(np.zeros(S, np.int32), np.full(S, pc.frames, np.uint32))
    dxid = dev.DeviceXidTable(layout, device)
    nf = S * (pc.frames + 1)
    dc_req = dev.decode_dev_autosize(ctx, d_req, nreq, ds_req, None, guess=(nf, 16, 2 * nf, nf // 4 + 16), device=device)
    dev.xidtab_build_dev(ctx, dc_req, S, dxid)
    dc_rsp = dev.decode_dev_autosize(ctx, d_rsp, nrsp, ds_rsp, dxid, guess=(nf, nf, 4 * nf, 16), device=device)
    torch.cuda.synchronize()

    counts = {}

    def step_dev(acc=None):
        rc = dev.decode_dev(ctx, d_req, nreq, ds_req, None, dc_req)
        assert rc == 0
        if acc is not None:
            acc.append(ctx.last_phase_ms())
        counts["slow_req"] = ctx.last_slow_frames
        dev.xidtab_build_dev(ctx, dc_req, S, dxid)
        rc = dev.decode_dev(ctx, d_rsp, nrsp, ds_rsp, dxid, dc_rsp)
        assert rc == 0
        if acc is not None:
            acc.append(ctx.last_phase_ms())
        counts["slow_rsp"] = ctx.last_slow_frames

    def barrier():
        if use_dist:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_dev()
    for name, dc in (("req", dc_req), ("rsp", dc_rsp)):
        d = dc.struct
        counts[name] = dict(frames=int(d.n_frames), stat=int(d.n_stat), buf=int(d.n_buf), acl=int(d.n_acl),
                            slow_frames=counts.pop("slow_" + name, None))
    frames_per_step = counts["req"]["frames"] + counts["rsp"]["frames"]

    import datetime
    sampler = ClockSampler(gpu_index_for(local_rank))
    sampler.start()
    for _ in range(2):          # keeps the GPU under the same load while the sampler comes up
        step_dev()
    phases = []
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
        fr = torch.tensor([frames_per_step, nreq + nrsp], device=device, dtype=torch.float64)
        dist.all_reduce(fr, op=dist.ReduceOp.SUM)
        total_frames, total_bytes = float(fr[0].item()), float(fr[1].item())
    else:
        total_frames, total_bytes = float(frames_per_step), float(nreq + nrsp)
    ms_per_step = ms / args.steps
    value = total_frames / (ms_per_step * 1e-3)

    # ---- roofline of the dominant kernel (k_decode), from the library's own CUDA events ----
    def out_bytes(c, with_emit):
        b = (c["frames"] * ROW_BYTES["frames_decode"] + c["stat"] * ROW_BYTES["stat"] +
             c["buf"] * ROW_BYTES["buf"] + c["acl"] * ROW_BYTES["acl"])
        return b + (c["frames"] * ROW_BYTES["frames_emit"] if with_emit else 0)

    alg = [nreq + out_bytes(counts["req"], False), nrsp + out_bytes(counts["rsp"], False)]
    dec_ms = [sum(p[2] for p in phases[0::2]) / max(1, len(phases[0::2])),
              sum(p[2] for p in phases[1::2]) / max(1, len(phases[1::2]))]
    split_ms = sum(p[0] + p[1] for p in phases) / max(1, args.steps)
    peak, peak_src = measured_peak()
    ach = (alg[0] + alg[1]) / ((dec_ms[0] + dec_ms[1]) * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "k_decode_traffic.json")
    if os.path.exists(tp):
        try:
            with open(tp) as f:
                tj = json.load(f)
            if tj.get("sessions") == S and tj.get("frames_per_session") == pc.frames:
                traffic = tj.get("dram_bytes_per_launch")
        except Exception:
            pass
    roofline = {
        "bound": "hbm", "kernel": "k_decode", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
        "peak_source": peak_src + " (MEASURED_PEAKS.json hbm_gbs)" if peak_src == "measured" else "fallback 6650",
        "frac_of_nominal_8TBps": ach / 8000.0,
        "traffic": traffic,
        "algorithmic_bytes_per_launch": (alg[0] + alg[1]) / 2,
        "launch_ms": {"request_streams": dec_ms[0], "response_streams": dec_ms[1]},
        "share_of_step": (dec_ms[0] + dec_ms[1]) / ms_per_step if not use_dist else None,
        "split_emit_ms_per_step": split_ms,
        "pipeline_GBps": (nreq + nrsp + out_bytes(counts["req"], True) + out_bytes(counts["rsp"], True))
        / (ms_per_step * 1e-3) / 1e9 if not use_dist else None,
    }

    # ---- end to end through the host-buffer entry points ----
    e2e_steps = max(2, min(args.steps, 4))
    h2d = d2h = 0

    def step_e2e():
        nonlocal h2d, d2h
        tot = {"frames": 0, "d2h": 0, "ok": 0}

        def on_slice(first, ns, rq, rs):
            # the "result read": every slice's columns are on the host; touch the status column
            tot["frames"] += rq.n_frames + rs.n_frames
            tot["d2h"] += rq.nbytes_columns() + rs.nbytes_columns()
            tot["ok"] += int((rs.meta & 0xff == 0).sum())

        ctx.replay_host(pc.req, pc.req_off, pc.rsp, pc.rsp_off, layout.base, layout.count, on_slice,
                        req_flags=pc.req_flags, rsp_flags=pc.rsp_flags, copy=False)
        h2d = nreq + nrsp + 2 * (S + 1) * 8 + 2 * S + 16 * S
        d2h = tot["d2h"]
        return tot["frames"]

    step_e2e()
    barrier()
    e0.record(ext)
    t_wall = time.perf_counter()
    got = 0
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

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "cfg2 mixed request/response replay, xid->opcode pairing",
                   "sessions_per_gpu": S, "frames_per_session": pc.frames,
                   "wire_bytes_per_gpu": nreq + nrsp, "frames_per_gpu": frames_per_step,
                   "l2": "inputs (%.1f GB per GPU) far larger than the 126 MB L2; no flush needed" % ((nreq + nrsp) / 1e9),
                   "parallelism": "sessions sharded over %d GPU(s), no collective" % world,
                   "seed": args.seed, "corpus_gen_s": round(gen_s, 1)},
        "wire_GBps": total_bytes / (ms_per_step * 1e-3) / 1e9,
        "rows_per_gpu": counts,
        "roofline": roofline,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps,
                "wire_GBps": total_bytes / (e2e_ms / e2e_steps * 1e-3) / 1e9,
                "pcie_roofline": pcie_roofline(total_bytes / (e2e_ms / e2e_steps * 1e-3) / 1e9 / max(world, 1))},
        "gpu_launches": int(launches),
        "split_fixups": ctx.split_fixups,
        "clocks": clocks,
    }

    # ---- the encode half of the path (rank 0, N=1 only): requests of the first sessions re-encoded
    # from their decoded columns, descriptors and arena resident in HBM ----
    if rank == 0 and world == 1 and not args.no_encode:
        ns_enc = min(S, 160)
        cut = int(pc.req_off[ns_enc])
        sub = ctx.decode_host(pc.req[:cut], pc.req_off[:ns_enc + 1], pc.req_flags[:ns_enc], None, None)
        rb, _ = zk.RequestBatch.from_columns(sub, pc.req[:cut])
        dr = dev.DeviceRequests(rb, device)
        out = None
        for _ in range(3):
            o0, o1, o2, enc_n, rc = dev.encode_dev(ctx, dr, cut, device=device, out=out)
            out = (o0, o1, o2)
            assert rc == 0
        l0 = ctx.kernel_launches
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        with torch.cuda.stream(ext):
            ev0.record()
            for _ in range(5):
                dev.encode_dev(ctx, dr, cut, device=device, out=out)
            ev1.record()
        torch.cuda.synchronize()
        enc_ms = ev0.elapsed_time(ev1) / 5
        payload = sum(int(np.maximum(rb.cols[k], 0).sum()) for k in ("path_len", "data_len", "elem_len",
                                                                    "acl_scheme_len", "acl_id_len"))
        alg = dr.nbytes() - int(rb.arena.size) - 32 + payload + enc_n + 9 * rb.n
        line["encode"] = {
            "packets": rb.n, "frame_bytes": enc_n, "ms": enc_ms, "packets_per_s": rb.n / (enc_ms * 1e-3),
            "frames_GBps": enc_n / (enc_ms * 1e-3) / 1e9,
            "roofline": {"bound": "hbm", "achieved": alg / (enc_ms * 1e-3) / 1e9, "peak": measured_peak()[0], "unit": "GB/s",
                         "frac": alg / (enc_ms * 1e-3) / 1e9 / measured_peak()[0],
                         "algorithmic_bytes": "descriptor columns + payload bytes read + frames, frame_off and status written"},
            "launches_per_call": (ctx.kernel_launches - l0) / 5,
            "sample": "requests of the first %d sessions, re-encoded from their decoded columns (zkb_encode_dev)" % ns_enc,
        }
        del dr, out

    # ---- CPU baseline beside it (rank 0, N=1 only) ----
    if rank == 0 and world == 1 and not args.no_cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import zkoracle as oracle
        threads = os.cpu_count() or 1
        rate, ns, f, nb, t = cpu_decode_rate(oracle, zk, pc, S, 1, threads, 8.0)
        rate_ip, ns2, f2, nb2, t2 = cpu_decode_rate(oracle, zk, pc, S, 0, threads, 4.0)
        rate_1, ns3, f3, nb3, t3 = cpu_decode_rate(oracle, zk, pc, S, 1, 1, 3.0)
        line["cpu_baseline"] = {
            "value": rate, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": "first %d of %d sessions, both directions (%.2f GB, %d frames, %.1f s): C port of the "
                      "reference loop, faithful mode (64 KiB chunks, per-frame copy + memmove)" % (ns, S, nb / 1e9, f, t),
            "in_place_value": rate_ip, "in_place_sample": "%d sessions, %.1f s, no per-frame copies" % (ns2, t2),
            "single_thread_value": rate_1,
            "node": "unavailable on this box (no node binary, no network)",
        }
    if rank == 0:
        print(json.dumps(line), flush=True)
    if use_dist:
        dist.destroy_process_group()
    ctx.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--gb", type=float, default=10.0, help="wire bytes per GPU, in GB")
    ap.add_argument("--seed", type=int, default=20240607)
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-encode", action="store_true", help="skip the encode leg")
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
