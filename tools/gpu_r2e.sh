#!/bin/bash
TAG=${1:-r02e}; O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -9 $O/${TAG}_pytest.log
for W in C4 C2; do
  R=""; [ $W = C4 ] && R="--reads 2000000"
  for CI in 1 0; do
  VGA_CHAIN_INT=$CI timeout 300 python bench.py --workload $W $R --steps 3 --warmup 2 --no-e2e --no-gaf --no-cpu-baseline > $O/${TAG}_${W}_ci$CI.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("$O/${TAG}_${W}_ci$CI.json"))
print("$W chain_int=$CI", {k: round(v,1) for k,v in d["stage_ms"].items()}, "value", round(d["value"]), "frac %.4f" % d["roofline_poa"]["frac"])
PY
  done
done
