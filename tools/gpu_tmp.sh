#!/bin/bash
O=gpurun_out; mkdir -p $O; TAG=${1:-r03c}
timeout 900 ncu --set full --clock-control none --import-source on -k regex:poa_fill_pair -c 1 -o $O/${TAG}_full_pair_C4 -f \
  python bench.py --steps 1 --warmup 0 --workload C4 --reads 250000 --batch-reads 250000 --no-cpu-baseline --no-e2e --no-gaf --no-index-build > $O/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
