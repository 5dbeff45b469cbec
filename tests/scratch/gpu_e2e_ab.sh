#!/bin/bash
# A/B of host-tick variants inside one box, interleaved
run() { env "$@" timeout 200 python bench.py --steps 40 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['e2e']['value']/1e6,2))"; }
for rep in 1 2 3; do
  echo "rep $rep: chunk launches $(run BLFIK_STREAMING=0) | streaming $(run BLFIK_STREAMING=1) | streaming memcpy-flags $(run BLFIK_WRITE_VALUE=0) | zero-copy $(run BLFIK_STREAMING=1 X=1 | tail -1)"
done
nvidia-smi --query-gpu=pcie.link.gen.current,pcie.link.width.current --format=csv,noheader
python tests/scratch/gpu_pcie.py | head -2
