#!/bin/bash
# bench variants given as env assignments. usage: tools/gpu_ab3.sh "<workloads>" VAR=1 VAR=2 ...
O=gpurun_out; mkdir -p $O
WL=$1; shift
for W in $WL; do for V in "$@"; do
env $V VGA_HOST_TIMELINE=1 timeout 300 python bench.py --steps 3 --warmup 2 --workload $W --no-cpu-baseline --no-e2e 2> $O/x.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$W $V', round(d['value']), {k:round(v,2) for k,v in d['stage_ms'].items()}, 'gcups', round(d['roofline_poa']['achieved'],1), 'launches', d['roofline_poa']['launches_per_step'])
"; done; done
