#!/bin/bash
# timeline of the POA launches + ncu --set full of the long-read fill kernel. usage: tools/gpu_prof_long.sh <tag> <workload> [env...]
TAG=$1; W=$2; shift 2
O=gpurun_out; mkdir -p $O
env "$@" VGA_HOST_TIMELINE=1 timeout 300 python bench.py --steps 1 --warmup 1 --workload $W --no-cpu-baseline --no-e2e 2>&1 | grep -E "poa (fill|trace)" | tail -12
env "$@" timeout 900 ncu --set full --clock-control none --import-source on -k regex:poa_fill_long -c 1 -o $O/${TAG}_full_long_$W -f \
  python bench.py --steps 1 --warmup 0 --workload $W --no-cpu-baseline --no-e2e > $O/${TAG}_ncu_long.log 2>&1; echo "ncu rc=$?"
