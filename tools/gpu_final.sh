#!/bin/bash
# last visit of a round: parity tests, smoke, full bench lines of the headline workloads, launch list. usage: tools/gpu_final.sh <tag>
TAG=$1; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py --steps 5 --warmup 3 > $O/${TAG}_bench_C2.json 2> $O/${TAG}_bench_C2.err; echo "bench C2 rc=$?"
for W in C4 C5; do
  timeout 400 python bench.py --steps 3 --warmup 3 --workload $W --cpu-seconds 6 > $O/${TAG}_bench_$W.json 2> $O/${TAG}_bench_$W.err; echo "bench $W rc=$?"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches_C2.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:chain_kernel -c 2 -o $O/${TAG}_full_chain_C2 -f \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $O/${TAG}_ncu_full_chain.log 2>&1; echo "ncu chain rc=$?"
for W in C2 C4 C5; do python -c "
import json; d=json.load(open('$O/${TAG}_bench_$W.json')); print('$W', round(d['value']), 'e2e', round(d['e2e']['value']), d['ms_steps'], d['e2e']['ms_steps'], {k:round(v,2) for k,v in d['stage_ms'].items()})"; done
