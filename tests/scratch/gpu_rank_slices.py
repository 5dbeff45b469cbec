"""Weak scaling: solve-kernel time of every rank's slice of the global batch (world size argv[1]) on one GPU, with the iteration
counts and hardness scores of its slowest robots."""
import sys; sys.path.insert(0, '/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
w = W.config(3); per = 65536; m = w.model
glob = W.synth_inputs(w, per * world, 3)
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
for rank in range(world):
    lo_, hi_ = W.shard_range(per * world, rank, world)
    inp = glob.slice(lo_, hi_)
    prob = capi.Problem(capi.Model(w.model), w.tasks, per); prob.set_inputs(inp); prob.kernel_timing(True)
    ks = []
    for rep in range(6):
        prob.set_state(inp.base_pos, inp.base_quat, inp.s); prob.sync(); flush.fill_(rep); torch.cuda.synchronize()
        prob.tick(1, w.dt); prob.sync(); ks.append(prob.kernel_times())
    st, it = prob.get_status()
    hi = np.minimum(0.5 * (m.q_upper - inp.s) / w.dt, 2.0); lo = np.maximum(0.5 * (m.q_lower - inp.s) / w.dt, -2.0)
    width = hi - lo
    score = (width.max(1, keepdims=True) / np.maximum(width, 1e-3) - 1).sum(1)
    heavy = np.flatnonzero(it >= 150)
    k = np.median(np.array(ks[2:]), axis=0)
    print(f"world {world} rank {rank}: assemble {k[0]:.3f} solve {k[1]:.3f} ms; unsolved {(st != 0).sum()}; robots >= 150 mat-vecs: iters {it[heavy].tolist()} scores {np.round(score[heavy], 2).tolist()} (share of robots with score 0: {(score == 0).mean():.2f})")
    prob.close()
