#!/bin/bash
O=gpurun_out; mkdir -p $O; TAG=${1:-r02s}
timeout 900 python -m pytest tests -m gpu -q -x -k "c4_real or synthetic_short or hla or pool or c2_full or c3_full" > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
for W in C4 C2; do
  R=""; [ $W = C4 ] && R="--reads 2000000"
  timeout 300 python bench.py --workload $W $R --steps 3 --warmup 2 --no-e2e --no-gaf --no-cpu-baseline --no-index-build > $O/${TAG}_$W.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("$O/${TAG}_$W.json"))
print("$W", {k: round(v,1) for k,v in d["stage_ms"].items()}, d["ms_steps"], "value", round(d["value"]))
PY
done
VGA_HOST_TIMELINE=1 timeout 300 python bench.py --workload C4 --reads 2000000 --steps 1 --warmup 2 --no-e2e --no-gaf --no-cpu-baseline --no-index-build 2>&1 >/dev/null | grep "^\[vga" | tail -24 > $O/${TAG}_timeline_C4.txt
tail -24 $O/${TAG}_timeline_C4.txt
