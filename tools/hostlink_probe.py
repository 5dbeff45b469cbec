#!/usr/bin/env python
"""Host-link ceiling of one box: every rank copies a pinned host buffer to its GPU (one plain cudaMemcpyAsync per buffer) and a
result buffer back, all ranks at once; prints the aggregate and per-GPU rates.  The e2e figure of bench.py at N GPUs is reported
as a fraction of what this probe allows.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/hostlink_probe.py
"""
import json
import os

import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    h2d_bytes, d2h_bytes = 65536 * 408, 65536 * 236          # the compact C3 records without the shared postural target; nu + status
    h_in = torch.empty(h2d_bytes, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(d2h_bytes, dtype=torch.uint8).pin_memory()
    d_in = torch.empty(h2d_bytes, dtype=torch.uint8, device="cuda")
    d_out = torch.empty(d2h_bytes, dtype=torch.uint8, device="cuda")
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    res = {}
    for name, both in (("h2d_only", False), ("h2d_and_d2h", True)):
        for rep in range(3):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s_in)
            for _ in range(20):
                with torch.cuda.stream(s_in):
                    d_in.copy_(h_in, non_blocking=True)
                if both:
                    with torch.cuda.stream(s_out):
                        h_out.copy_(d_out, non_blocking=True)
            s_in.wait_stream(s_out)
            e1.record(s_in)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 20
        t = torch.tensor([ms], device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res[name] = {"ms_per_step_max_over_ranks": float(t[0]), "h2d_GBps_per_gpu": h2d_bytes / float(t[0]) * 1e-6,
                     "h2d_GBps_aggregate": world * h2d_bytes / float(t[0]) * 1e-6,
                     "ticks_per_s_ceiling": world * 65536 / (float(t[0]) * 1e-3)}
    if rank == 0:
        print(json.dumps({"n_gpus": world, "h2d_bytes_per_gpu": h2d_bytes, "d2h_bytes_per_gpu": d2h_bytes, **res}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
