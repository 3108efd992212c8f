#!/bin/bash
TAG=${1:-r02g}; O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x --durations=3 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/${TAG}_pytest.log
for W in C4 C2 C5; do
  R=""; [ $W = C4 ] && R="--reads 2000000"; [ $W = C5 ] && R="--reads 25000"
  for CT in 1 0; do
  VGA_TABLE_COMPACT=$CT timeout 300 python bench.py --workload $W $R --steps 3 --warmup 2 --no-e2e --no-gaf --no-cpu-baseline > $O/${TAG}_${W}_ct$CT.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("$O/${TAG}_${W}_ct$CT.json"))
print("$W compact=$CT", {k: round(v,2) for k,v in d["stage_ms"].items()}, "value", round(d["value"]), "lookup frac %.3f" % d["roofline_lookup"]["frac"], "image MB", d["setup"]["index_image_bytes"]>>20)
PY
  done
done
