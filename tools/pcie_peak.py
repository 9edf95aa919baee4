"""Development aid: pinned host <-> device copy bandwidth of this box with N GPUs copying at the
same time (the PCIe roofline of the end-to-end numbers).

    python tools/pcie_peak.py                       # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29611 tools/pcie_peak.py      # N GPUs at once, one process each

Every rank copies 2 GiB pinned -> device, device -> pinned, and both at once in the replay's byte
mix (2 GiB up while 1 GiB comes down); the ranks start each test together (gloo barrier) and the
slowest rank's time counts.  Rank 0 prints one JSON object: per-GPU GB/s for each test.
"""
import json
import os
import time

import torch

rank = int(os.environ.get("RANK", 0))
world = int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
if world > 1:
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("gloo")
torch.cuda.set_device(local)
n = 2 << 30
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(n // 2, dtype=torch.uint8, pin_memory=True)
d2 = torch.empty(n // 2, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    if world > 1:
        dist.barrier()


def timed(fn, reps=3):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        barrier()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        best = min(best, dt)
    return best


def both():
    with torch.cuda.stream(s1):
        d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)


a = timed(lambda: d.copy_(h, non_blocking=True))
b = timed(lambda: h.copy_(d, non_blocking=True))
c = timed(both)
if rank == 0:
    numa = None
    try:
        numa = len([x for x in os.listdir("/sys/devices/system/node") if x.startswith("node")])
    except Exception:
        pass
    print(json.dumps({
        "how": "tools/pcie_peak.py, %d GPU(s) copying at once: 2 GiB pinned<->device per GPU, best of 3, slowest rank" % world,
        "n_gpus": world, "h2d_gbs": n / a / 1e9, "d2h_gbs": n / b / 1e9, "duplex_h2d_equiv_gbs": n / c / 1e9,
        "aggregate_h2d_gbs": world * n / a / 1e9, "aggregate_duplex_h2d_equiv_gbs": world * n / c / 1e9,
        "host_cpus": os.cpu_count(), "numa_nodes": numa,
        "note": "duplex = 2 GiB H2D concurrent with 1 GiB D2H per GPU (the byte mix of the cfg2 replay); "
                "time expressed as H2D bytes per second per GPU"}))
if world > 1:
    dist.destroy_process_group()
