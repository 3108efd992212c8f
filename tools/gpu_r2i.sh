#!/bin/bash
TAG=${1:-r02i}; O=gpurun_out; mkdir -p $O
for MB in 4 5 6; do
  VGA_POA_MINB=$MB timeout 300 python bench.py --workload C4 --reads 2000000 --steps 3 --warmup 2 --no-e2e --no-gaf --no-cpu-baseline > $O/${TAG}_C4_mb$MB.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("$O/${TAG}_C4_mb$MB.json"))
print("C4 minb=$MB", "ms_poa %.1f ms_trace %.1f total %.1f" % (d["stage_ms"]["ms_poa"], d["stage_ms"]["ms_trace"], d["stage_ms"]["ms_total"]))
PY
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:poa_fill_pair -c 1 -o $O/${TAG}_full_pair_C4 -f \
  python bench.py --steps 1 --warmup 0 --workload C4 --reads 1000000 --no-cpu-baseline --no-e2e --no-gaf > $O/${TAG}_ncu_pair.log 2>&1; echo "ncu pair rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:poa_trace_planes -c 1 -o $O/${TAG}_full_tracep_C4 -f \
  python bench.py --steps 1 --warmup 0 --workload C4 --reads 1000000 --no-cpu-baseline --no-e2e --no-gaf > $O/${TAG}_ncu_tracep.log 2>&1; echo "ncu trace rc=$?"
ls -la $O/*.ncu-rep
