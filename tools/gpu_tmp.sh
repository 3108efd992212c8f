#!/bin/bash
O=gpurun_out; mkdir -p $O; TAG=${1:-r03a}
timeout 900 python -m pytest tests -m gpu -q -x -k "c1 or hla or synthetic or repeats or c2_full or c4_real or reverse_strand or chains_only or medium" > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
for W in C4 C3b C2; do
  R=""; [ $W = C4 ] && R="--reads 2000000"; [ $W = C3b ] && R="--reads 250000"
  timeout 300 python bench.py --workload $W $R --steps 3 --warmup 2 --no-e2e --no-gaf --no-cpu-baseline --no-index-build > $O/${TAG}_$W.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("$O/${TAG}_$W.json"))
print("$W", {k: round(v,1) for k,v in d["stage_ms"].items()}, "value", round(d["value"]))
PY
done
