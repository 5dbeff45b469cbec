"""Generates tests/golden/*.npz from the CPU oracle.

The reference holds NO golden numeric vectors for this path and cannot be built or imported in this
image, so these fixtures are *not* reference outputs: they freeze the oracle's results on
seeded inputs (after the oracle itself passed its independent checks in tests/test_oracle_*.py) so that
(a) a later change of the oracle is noticed and (b) the GPU box, which has no /root/reference either,
compares the CUDA path against committed numbers as well as against the live oracle.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import blf_b200  # noqa: E402,F401
from blf_b200 import workloads as W  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
N = 16


def main():
    for cid in (1, 3, 4):
        w = W.config(cid)
        inp = W.synth_inputs(w, N, cid)
        o = Oracle(w.model, w.tasks)
        frames = sorted({t.frame for t in w.tasks if t.frame >= 0})
        fp, fR, fJ, com, cJ, Hs, gs = [], [], [], [], [], [], []
        for i in range(N):
            R, P = o.fk(inp.base_pos[i], inp.base_quat[i], inp.s[i])
            poses = [o.frame_pose(R, P, f) for f in frames]
            fp.append([x[1] for x in poses]); fR.append([x[0] for x in poses])
            fJ.append([o.frame_jacobian(R, P, f) for f in frames])
            c, J = o.com(R, P)
            com.append(c); cJ.append(J)
            H, g, C, l, u = o.assemble(inp.base_pos[i], inp.base_quat[i], inp.s[i], [None if sp is None else sp[i] for sp in inp.setpoints])
            Hs.append(H); gs.append(g)
        p, q, s = inp.base_pos.copy(), inp.base_quat.copy(), inp.s.copy()
        nu, status, _ = o.tick_batch(p, q, s, inp.setpoints, dt_integrate=w.dt)
        out = dict(base_pos=inp.base_pos, base_quat=inp.base_quat, s=inp.s, frames=np.array(frames), frame_pos=np.array(fp),
                   frame_rot=np.array(fR), frame_jac=np.array(fJ), com=np.array(com), com_jac=np.array(cJ), H=np.array(Hs), g=np.array(gs),
                   nu=nu, status=status, next_pos=p, next_quat=q, next_s=s)
        for t, sp in enumerate(inp.setpoints):
            if sp is not None:
                out[f"sp{t}"] = sp
        np.savez_compressed(os.path.join(HERE, f"golden_c{cid}.npz"), **out)
        print("wrote", f"golden_c{cid}.npz", "status", np.bincount(status))


if __name__ == "__main__":
    main()
