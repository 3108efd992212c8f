#!/bin/bash
# parity tests for both pair instances, then short benches with per-launch timings
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/ab2_pytest_g8.log 2>&1; echo "pytest(G8) rc=$?"; tail -12 $O/ab2_pytest_g8.log
VGA_POA_G4=1 timeout 900 python -m pytest tests -m gpu -x -q > $O/ab2_pytest_g4.log 2>&1; echo "pytest(G4) rc=$?"; tail -12 $O/ab2_pytest_g4.log
for W in ${1:-C2 C3a C3b C4}; do for V in VGA_POA_PAIR=1 VGA_POA_G4=1; do
env $V VGA_HOST_TIMELINE=1 timeout 300 python bench.py --steps 3 --warmup 2 --workload $W --no-cpu-baseline --no-e2e 2> $O/x.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$W $V', round(d['value']), {k:round(v,2) for k,v in d['stage_ms'].items()}, 'gcups', round(d['roofline_poa']['achieved'],1), 'launches', d['roofline_poa']['launches_per_step'])
"; grep -E "poa (fill|trace)" $O/x.err | tail -3; done; done
