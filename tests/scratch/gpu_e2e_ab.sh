#!/bin/bash
run() { env "$@" timeout 200 python bench.py --steps 40 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['e2e']['value']/1e6,2))"; }
for rep in 1 2; do
  echo "rep $rep: ramp1.5 p12 $(run BLFIK_PIECES=12 BLFIK_RAMP=1) | smooth p8 $(run BLFIK_PIECES=8) | smooth p12 $(run BLFIK_PIECES=12) | smooth p16 $(run BLFIK_PIECES=16) | smooth p20 $(run BLFIK_PIECES=20) | smooth p24 $(run BLFIK_PIECES=24)"
done
