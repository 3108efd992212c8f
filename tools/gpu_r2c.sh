#!/bin/bash
TAG=${1:-r02c}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x -k "wide or c5 or 1024 or long" --durations=8 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $O/${TAG}_pytest.log
timeout 600 python bench.py --workload C5 --steps 2 --warmup 1 --e2e-steps 1 --no-gaf --cpu-seconds 6 > $O/${TAG}_bench_C5.json 2> $O/${TAG}_bench_C5.err; echo "bench rc=$?"; tail -3 $O/${TAG}_bench_C5.err
python - <<PY
import json
d=json.load(open("$O/${TAG}_bench_C5.json"))
print("C5 value", round(d["value"]), "e2e", round(d["e2e"]["value"]), {k: round(v,1) for k,v in d["stage_ms"].items()}, d["roofline_poa"]["frac"], d.get("parity_sample"))
PY
