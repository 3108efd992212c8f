#!/bin/bash
O=gpurun_out; mkdir -p $O; TAG=${1:-r02m}
timeout 1500 python -m pytest tests -m gpu -q -x --durations=3 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/${TAG}_pytest.log
for E in 1 0; do
echo "== VGA_POA_EARLY=$E"
VGA_POA_EARLY=$E VGA_HOST_TIMELINE=1 timeout 300 python bench.py --workload C3a --reads 125000 --steps 2 --warmup 1 --no-e2e --no-gaf --no-cpu-baseline 2> $O/${TAG}_tl_$E.txt > $O/${TAG}_C3a_e$E.json
grep -E "poa (fill|trace)" $O/${TAG}_tl_$E.txt | tail -4
python - <<PY
import json
d=json.load(open("$O/${TAG}_C3a_e$E.json"))
print("C3a early=$E", {k: round(v,2) for k,v in d["stage_ms"].items()}, "value", round(d["value"]), "poa frac %.4f" % d["roofline_poa"]["frac"])
PY
done
