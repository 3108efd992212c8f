#!/bin/bash
O=gpurun_out; mkdir -p $O
for CHK in 1 0; do
echo "== VGA_POA_CHK=$CHK"
VGA_POA_CHK=$CHK VGA_HOST_TIMELINE=1 timeout 300 python bench.py --workload C3a --reads 125000 --steps 1 --warmup 1 --no-e2e --no-gaf --no-cpu-baseline 2>&1 >/dev/null | grep -E "poa (fill|trace)" | tail -8
done
