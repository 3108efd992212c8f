#!/bin/bash
# parity tests, then long runs that print the host timeline only of slow calls (hunting sporadic slow steps). usage: tools/gpu_ab8.sh <tag>
TAG=$1; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/${TAG}_pytest.log
env VGA_HOST_TIMELINE=33 timeout 300 python bench.py --steps 40 --warmup 3 --workload C2 --no-cpu-baseline 2> $O/${TAG}_C2.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('C2', round(d['value']), round(d['ms_per_step'],2), d['ms_steps'], 'e2e', d['e2e']['ms_steps'])"
grep "^\[vga" $O/${TAG}_C2.err | head -80
for W in C4 C5; do
timeout 300 python bench.py --steps 5 --warmup 3 --workload $W --no-cpu-baseline 2> $O/${TAG}_$W.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$W', round(d['value']), round(d['ms_per_step'],2), d['ms_steps'], 'e2e', d['e2e']['ms_steps'], {k:round(v,2) for k,v in d['stage_ms'].items()})"
done
env VGA_HOST_TIMELINE=1 timeout 300 python bench.py --steps 1 --warmup 2 --workload C4 --no-cpu-baseline --no-e2e 2>&1 >/dev/null | grep "^\[vga-tl" | tail -14
