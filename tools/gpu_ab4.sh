#!/bin/bash
# parity tests + C2/C4 bench lines + ncu --set full of the traceback kernel. usage: tools/gpu_ab4.sh <tag>
TAG=$1; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/${TAG}_pytest.log
for W in C2 C4; do
  timeout 300 python bench.py --steps 5 --warmup 3 --workload $W --no-cpu-baseline --no-e2e 2> $O/${TAG}_$W.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$W', round(d['value']), d['ms_per_step'], d['stage_ms'])"
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:poa_trace -c 1 -o $O/${TAG}_full_trace_C2 -f \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $O/${TAG}_ncu_trace.log 2>&1; echo "ncu trace rc=$?"
