import sys, time; sys.path.insert(0,'/root/repo')
import numpy as np
import blf_b200
from blf_b200 import workloads as W
from oracle.oracle import Oracle
w=W.config(3); N=65536; inp=W.synth_inputs(w,N,3); o=Oracle(w.model,w.tasks)
t0=time.time()
nu,st,it=o.tick_batch(inp.base_pos.copy(),inp.base_quat.copy(),inp.s.copy(),inp.setpoints,nthreads=8)
print('oracle time',time.time()-t0,'status',np.bincount(st),'hard idx',np.nonzero(st)[0], 'iters of hard', it[st!=0], 'iters pct', np.percentile(it,[50,90,99,99.9]), it.max())
np.save('/tmp/oracle_c3_status.npy', st); np.save('/tmp/oracle_c3_iters.npy', it); np.save('/tmp/oracle_c3_nu.npy', nu)
