#!/bin/bash
O=gpurun_out; mkdir -p $O; TAG=${1:-r02r}
for MB in 4 5 6; do
  VGA_POA_MINB=$MB timeout 300 python -m pytest tests -m gpu -q -x -k "c4_real or synthetic_short or hla" > $O/${TAG}_pytest_mb$MB.log 2>&1; echo "pytest minb=$MB rc=$?"
  VGA_POA_MINB=$MB timeout 300 python bench.py --workload C4 --reads 2000000 --steps 3 --warmup 2 --no-e2e --no-gaf --no-cpu-baseline --no-index-build > $O/${TAG}_C4_mb$MB.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("$O/${TAG}_C4_mb$MB.json"))
print("C4 minb=$MB", "ms_poa %.1f ms_trace %.1f total %.1f" % (d["stage_ms"]["ms_poa"], d["stage_ms"]["ms_trace"], d["stage_ms"]["ms_total"]), d["ms_steps"])
PY
done
VGA_HOST_TIMELINE=1 timeout 300 python bench.py --workload C4 --reads 2000000 --steps 1 --warmup 2 --no-e2e --no-gaf --no-cpu-baseline --no-index-build 2>&1 >/dev/null | grep "^\[vga" | tail -40 > $O/${TAG}_timeline_C4.txt
tail -40 $O/${TAG}_timeline_C4.txt
