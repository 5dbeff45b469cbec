"""world_size-2 gloo test of the N>1 host-side path (sharding, gather, max-over-ranks) on the CPU.
The per-shard compute stand-in here is the CPU oracle (the product's compute needs a GPU)."""
import os
import socket
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_global, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import blf_b200  # noqa: F401
    from blf_b200 import sharding, workloads as W
    from oracle.oracle import Oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    w = W.config(3)
    a, b = sharding.shard_range(n_global, rank, world)
    inp = W.synth_inputs(w, n_global, 3).slice(a, b)
    nu, status, _ = Oracle(w.model, w.tasks).tick_batch(inp.base_pos, inp.base_quat, inp.s, inp.setpoints)
    full = sharding.gather_outputs(nu, n_global, rank, world)
    tmax = sharding.max_over_ranks([1.0 + rank, 5.0 - rank], world)
    tsum = sharding.sum_over_ranks([float(b - a)], world)
    dist.barrier()
    if rank == 0:
        q.put((full, tmax, tsum))
    dist.destroy_process_group()


def test_two_rank_gloo_gather_matches_single_process():
    import torch.multiprocessing as mp
    n_global = 21                      # odd: ragged 10 / 11 split
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_global, q)) for r in range(2)]
    for p in procs:
        p.start()
    full, tmax, tsum = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    sys.path.insert(0, ROOT)
    import blf_b200  # noqa: F401
    from blf_b200 import workloads as W
    from oracle.oracle import Oracle
    w = W.config(3)
    inp = W.synth_inputs(w, n_global, 3)
    ref, _, _ = Oracle(w.model, w.tasks).tick_batch(inp.base_pos, inp.base_quat, inp.s, inp.setpoints)
    assert full.shape == ref.shape and np.array_equal(full, ref)
    assert tmax == [2.0, 5.0] and tsum == [float(n_global)]
