"""CPU, world_size 2 over gloo: the N>1 host path (bundle dealing + ordered merge, bramble_b200/shard.py).
Each rank projects its own bundles (with the CPU oracle standing in for a GPU here -- test only) and rank 0
must end up with exactly the single-process result in bundle order."""
import os
import socket
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from bramble_b200 import shard
    import common
    import oracle_lib
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    wl = common.make_workload("default", 77, n_names=900, n_genes=20, chunk_size=300)
    nb = len(wl["batches"])
    mine = {k: oracle_lib.port_project(wl["index"], wl["params"], wl["batches"][k]) for k in shard.bundles_of_rank(nb, world, rank)}
    merged = shard.gather_in_order(mine, nb, world, rank, dist)
    if rank == 0:
        single = [oracle_lib.port_project(wl["index"], wl["params"], b) for b in wl["batches"]]
        ok = len(merged) == nb and nb >= 3
        for a, b in zip(merged, single):
            ok = ok and a["n_rec"] == b["n_rec"] and np.array_equal(a["tid"], b["tid"]) and np.array_equal(a["pos"], b["pos"])
        q.put(bool(ok))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_merge_in_bundle_order():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
    assert ok
