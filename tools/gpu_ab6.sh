#!/bin/bash
# parity tests, bench stage times for the given workloads, host timeline of one C2 call. usage: tools/gpu_ab6.sh <tag> "<workloads>"
TAG=$1; WL=$2; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/${TAG}_pytest.log
for W in $WL; do
  timeout 300 python bench.py --steps 4 --warmup 3 --workload $W --no-cpu-baseline --no-e2e 2> $O/${TAG}_${W}.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$W', round(d['value']), round(d['ms_per_step'],2), d['ms_steps'], {k:round(v,2) for k,v in d['stage_ms'].items()})"
done
env VGA_HOST_TIMELINE=1 timeout 300 python bench.py --steps 1 --warmup 2 --workload C2 --no-cpu-baseline --no-e2e 2>&1 >/dev/null | grep -v "^\[vga\] poa" | tail -45
