#!/bin/bash
# One GPU-box visit: parity tests, bench lines for every workload, ncu launch list and one --set full capture.
# usage: tools/gpu_round.sh <tag> [skip_tests]
TAG=${1:-rXX}
O=gpurun_out
mkdir -p $O
if [ -z "$2" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest.log
  tail -3 $O/${TAG}_pytest.log
fi
timeout 300 python bench.py --steps 5 --warmup 3 > $O/${TAG}_bench_C2.json 2> $O/${TAG}_bench_C2.err; echo "bench C2 rc=$?"
for W in C3a C3b C4 C5; do
  timeout 400 python bench.py --steps 3 --warmup 3 --workload $W --cpu-seconds 6 > $O/${TAG}_bench_$W.json 2> $O/${TAG}_bench_$W.err; echo "bench $W rc=$?"
done
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_ref_C2.json 2> $O/${TAG}_bench_ref_C2.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches_C2.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:poa_fill_pair_kernel -c 2 -o $O/${TAG}_full_poa_pair_C2 -f \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:seed_ -c 2 -o $O/${TAG}_full_seed_C4 -f \
  python bench.py --steps 1 --warmup 1 --workload C4 --reads 400000 --no-cpu-baseline --no-e2e > $O/${TAG}_ncu_full_seed.log 2>&1; echo "ncu seed rc=$?"
cat $O/${TAG}_bench_C2.json | head -c 3000
echo
timeout 900 ncu --set full --clock-control none --import-source on -k regex:poa_fill_long -c 1 -o $O/${TAG}_full_poa_long_C5 -f \
  python bench.py --steps 1 --warmup 0 --workload C5 --no-cpu-baseline --no-e2e > $O/${TAG}_ncu_full_long.log 2>&1; echo "ncu long rc=$?"
for W in C2 C4; do
  env VGA_HOST_TIMELINE=1 timeout 300 python bench.py --steps 1 --warmup 2 --workload $W --no-cpu-baseline --no-e2e 2>&1 >/dev/null | grep "^\[vga" | tail -22 > $O/${TAG}_timeline_$W.txt
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:chain_kernel -c 1 -o $O/${TAG}_full_chain_C2 -f \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $O/${TAG}_ncu_full_chain.log 2>&1; echo "ncu chain rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:poa_trace -c 1 -o $O/${TAG}_full_trace_C2 -f \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $O/${TAG}_ncu_full_trace.log 2>&1; echo "ncu trace rc=$?"
