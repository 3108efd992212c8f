#!/bin/bash
TAG=${1:-r02j}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x -k "c1 or hla or synthetic_short or edge or c2_full or batch_ind or size_ind or c4_real or compact" > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${TAG}_pytest.log
for W in C4 C2; do
  R=""; [ $W = C4 ] && R="--reads 2000000"
  timeout 300 python bench.py --workload $W $R --steps 3 --warmup 2 --no-e2e --no-gaf --no-cpu-baseline > $O/${TAG}_${W}.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("$O/${TAG}_${W}.json"))
print("$W", {k: round(v,2) for k,v in d["stage_ms"].items()}, "value", round(d["value"]), "poa frac %.4f" % d["roofline_poa"]["frac"])
PY
done
