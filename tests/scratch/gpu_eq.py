"""Equality-only kernel: time and iteration counts of C2 for a few proximal weights (sigma)."""
import sys; sys.path.insert(0, '/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
cid = int(sys.argv[1]) if len(sys.argv) > 1 else 2
N = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
w = W.config(cid); inp = W.synth_inputs(w, N, cid)
model = capi.Model(w.model); prob = capi.Problem(model, w.tasks, N)
prob.set_inputs(inp)
stream = torch.cuda.ExternalStream(prob.stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
ref = None
for sigma in ([float(a) for a in sys.argv[3:]] or [1e-2]):
    st = prob.get_solver_settings() if hasattr(prob, "get_solver_settings") else None
    prob.set_solver(sigma=sigma)
    ts = []
    for rep in range(8):
        prob.set_state(inp.base_pos, inp.base_quat, inp.s); prob.sync()
        flush.fill_(rep); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); prob.tick(1, w.dt); e1.record(stream); prob.sync(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    nu = prob.get_output(); stt, it = prob.get_status()
    if ref is None: ref = nu.copy()
    ts = np.array(ts[2:])
    print(f"C{cid} N={N} path {prob.kernel_path} sigma {sigma:g}: median {np.median(ts):.3f} ms {N / np.median(ts) / 1e3:.2f} Mticks/s  unsolved {int((stt != 0).sum())} "
          f"iters median {np.median(it):.0f} max {it.max()}  |nu - nu(first sigma)| {np.abs(nu - ref).max():.2e}")
