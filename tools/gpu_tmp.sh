#!/bin/bash
O=gpurun_out; mkdir -p $O; TAG=${1:-r02k}
B="--no-cpu-baseline --no-e2e --no-gaf --no-index-build"
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); s=d["stage_ms"]
    print(sys.argv[1].split('/')[-1], "value", round(d["value"]), "ms/step", round(d["ms_per_step"],1), {k[3:]:round(v,1) for k,v in s.items()})
except Exception as e: print(sys.argv[1], "ERR", e)
PY
}
timeout 1500 python -m pytest tests -m gpu -q -x > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${TAG}_pytest.log
timeout 300 python bench.py --steps 3 --warmup 2 --workload C4 --reads 2000000 $B > $O/${TAG}_c4.json 2>$O/${TAG}_c4.err; line $O/${TAG}_c4.json
timeout 300 python bench.py --steps 3 --warmup 2 --workload C2 $B > $O/${TAG}_c2.json 2>$O/${TAG}_c2.err; line $O/${TAG}_c2.json
timeout 300 python bench.py --steps 2 --warmup 2 --workload C5 $B > $O/${TAG}_c5.json 2>$O/${TAG}_c5.err; line $O/${TAG}_c5.json
timeout 300 python bench.py --steps 3 --warmup 2 --workload C3b $B > $O/${TAG}_c3b.json 2>$O/${TAG}_c3b.err; line $O/${TAG}_c3b.json
