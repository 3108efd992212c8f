#!/bin/bash
# A/B of the POA fill kernels: parity tests, then short bench runs per variant.  usage: tools/gpu_ab.sh <tag> "<workloads>"
TAG=${1:-ab}; WL=${2:-C2}
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_g8.log 2>&1; echo "pytest(G8) rc=$?"; tail -15 $O/${TAG}_pytest_g8.log
VGA_POA_G4=1 timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_g4.log 2>&1; echo "pytest(G4) rc=$?"; tail -15 $O/${TAG}_pytest_g4.log
for W in $WL; do
  for V in "VGA_POA_PAIR=0" "VGA_POA_PAIR=1" "VGA_POA_G4=1"; do
    env $V timeout 300 python bench.py --steps 3 --warmup 3 --workload $W --no-cpu-baseline --no-e2e > $O/${TAG}_${W}_${V}.json 2> $O/${TAG}_${W}_${V}.err
    echo "$W $V rc=$? $(python -c "
import json,sys
try:
    d=json.load(open('$O/${TAG}_${W}_${V}.json')); print(round(d['value']), {k:round(v,2) for k,v in d['stage_ms'].items()}, 'gcups', round(d['roofline_poa']['achieved'],1), 'launches', d['roofline_poa']['launches_per_step'])
except Exception as e: print('ERR', e)
")"
  done
done
