#!/bin/bash
# One GPU-box visit that produces everything profiles/ holds for a round.  usage: tools/gpu_evidence.sh <tag> <git sha>   (gpurun, 1 GPU)
#   parity tests + smoke, the default bench line (C4 10 M reads + C5 record), C2 / C3a / C3b lines, the reference arm,
#   ncu launch list of a short default bench, ncu --set full of the dominant kernels on the CURRENT build, host timelines.
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
echo "${2:-unknown}" > $O/${TAG}_sha.txt      # the box has no .git: the caller passes the SHA of the tree it sends
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 1200 python bench.py --steps 3 --warmup 3 > $O/${TAG}_bench_default.json 2> $O/${TAG}_bench_default.err; echo "bench default rc=$?"
for W in C2 C3a C3b; do
  timeout 400 python bench.py --steps 5 --warmup 3 --workload $W --cpu-seconds 6 --no-index-build > $O/${TAG}_bench_$W.json 2> $O/${TAG}_bench_$W.err; echo "bench $W rc=$?"
done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err; echo "ref rc=$?"
# ---- ncu: launch list of the bench command on a bounded workload (the share of each kernel in a step), full captures
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $O/${TAG}_launches_C4.csv \
  python bench.py --steps 1 --warmup 1 --workload C4 --reads 1000000 --no-cpu-baseline --no-e2e --no-gaf --no-index-build > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
cap() {  # cap <name> <kernel regex> <workload> <reads> [skip]
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$2 -s ${5:-0} -c 1 -o $O/${TAG}_full_$1 -f \
    python bench.py --steps 1 --warmup 0 --workload $3 --reads $4 --batch-reads $4 --no-cpu-baseline --no-e2e --no-gaf --no-index-build > $O/${TAG}_ncu_$1.log 2>&1; echo "ncu $1 rc=$?"
}
cap pair_C4 poa_fill_pair C4 250000
cap seed_count_C4 seed_count C4 1000000
cap seed_emit_C4 seed_emit C4 1000000
cap chain_C4 chain_kernel C4 250000
cap trace_C4 poa_trace C4 250000
cap long_C5 poa_fill_long C5 12500
cap wide_C5 poa_fill_wide C5 12500
cap pair_C2 poa_fill_pair C2 100000
# the cell / byte counts of exactly those launches (same --reads, one call): denominators of bytes per cell
for spec in "C4 250000" "C4 1000000" "C5 12500" "C2 100000"; do set -- $spec
  timeout 300 python bench.py --steps 1 --warmup 1 --workload $1 --reads $2 --batch-reads $2 --no-cpu-baseline --no-e2e --no-gaf --no-index-build > $O/${TAG}_counts_$1_$2.json 2>/dev/null
done
for W in C4 C5; do
  R="--reads 2000000"; [ $W = C5 ] && R="--reads 25000"
  env VGA_HOST_TIMELINE=1 timeout 300 python bench.py --steps 1 --warmup 2 --workload $W $R --no-cpu-baseline --no-e2e --no-gaf --no-index-build 2>&1 >/dev/null | grep "^\[vga" | tail -40 > $O/${TAG}_timeline_$W.txt
done
ls -la $O | grep ${TAG} | awk '{print $5, $9}'
