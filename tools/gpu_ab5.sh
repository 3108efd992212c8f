#!/bin/bash
# A/B of environment knobs on bench stage times. usage: tools/gpu_ab5.sh <tag> <workloads> <env1> <env2> ...   (env "X=1" = no-op label)
TAG=$1; WL=$2; shift 2; O=gpurun_out; mkdir -p $O
for W in $WL; do for E in "$@"; do
  env $E timeout 300 python bench.py --steps 4 --warmup 3 --workload $W --no-cpu-baseline --no-e2e 2> $O/${TAG}_${W}_$E.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$W $E', round(d['value']), round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['stage_ms'].items()})"
done; done
env VGA_HOST_TIMELINE=1 timeout 300 python bench.py --steps 1 --warmup 2 --workload C2 --no-cpu-baseline --no-e2e 2>&1 >/dev/null | grep "\[vga\]" | tail -40
