#!/usr/bin/env python
"""Host<->device copy ceiling of this box, per rank and with all ranks copying at once (run under torchrun).
The end-to-end figure of bench.py at N GPUs is bounded by these numbers: every step moves ~0.2 GB host->device and
~0.43 GB device->host per rank.  Usage:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/pcie_probe.py [--bind]
--bind pins each rank (and so its first-touched pinned pages) to the CPUs of the GPU's NUMA node first."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import torch.distributed as dist
from bramble_b200 import hostbind

ap = argparse.ArgumentParser(); ap.add_argument("--bind", action="store_true"); ap.add_argument("--mb", type=int, default=512)
args = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
info = hostbind.bind_to_gpu_node(local) if args.bind else hostbind.describe(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = args.mb << 20
h_in, h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True), torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_in.fill_(1); h_out.fill_(2)
d_in, d_out = torch.empty(n, dtype=torch.uint8, device="cuda"), torch.ones(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

def bar():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()

def run(h2d, d2h, reps=4):
    bar(); t0 = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    return reps * n / dt / 1e9

res = {"rank": rank, **info}
run(True, True, 1)
# alone: ranks take turns
for who in range(world):
    bar()
    if who == rank:
        torch.cuda.synchronize()
        t = []
        for h, d in ((1, 0), (0, 1), (1, 1)):
            t0 = time.perf_counter()
            for _ in range(4):
                if h:
                    with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
                if d:
                    with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
            torch.cuda.synchronize(); t.append(4 * n / (time.perf_counter() - t0) / 1e9)
        res["alone_h2d"], res["alone_d2h"], res["alone_both_each_dir"] = [round(x, 1) for x in t]
bar()
res["all_h2d"], res["all_d2h"], res["all_both_each_dir"] = round(run(1, 0), 1), round(run(0, 1), 1), round(run(1, 1), 1)
out = [None] * world
if world > 1:
    dist.all_gather_object(out, res)
else:
    out = [res]
if rank == 0:
    for r in out:
        print(json.dumps(r))
    print(json.dumps({"sum_all_h2d": round(sum(r["all_h2d"] for r in out), 1), "sum_all_d2h": round(sum(r["all_d2h"] for r in out), 1),
                      "sum_all_both_each_dir": round(sum(r["all_both_each_dir"] for r in out), 1), "host_cpus": os.cpu_count(),
                      "unit": "GB/s", "bind": args.bind}))
if world > 1:
    dist.destroy_process_group()
