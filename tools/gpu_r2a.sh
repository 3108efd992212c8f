#!/bin/bash
# round-2 visit A: parity tests (incl. the full-size configs), smoke, the new default bench line (C4 10 M + C5 100 k), C2 line.
TAG=${1:-r02a}; O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader; nproc; free -g | head -2
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 $O/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py --steps 3 --warmup 3 > $O/${TAG}_bench_default.json 2> $O/${TAG}_bench_default.err; echo "bench default rc=$?"; tail -5 $O/${TAG}_bench_default.err
timeout 300 python bench.py --steps 5 --warmup 3 --workload C2 > $O/${TAG}_bench_C2.json 2> $O/${TAG}_bench_C2.err; echo "bench C2 rc=$?"
python - <<PY
import json
for f in ("default","C2"):
    try:
        d=json.load(open("$O/${TAG}_bench_%s.json" % f))
    except Exception as e:
        print(f, "no json", e); continue
    for name, r in (("main", d), ("long", d.get("long_reads"))):
        if not r: continue
        print(f, name, "value", round(r["value"]), "e2e", round(r["e2e"]["value"]), "pageable", round(r["e2e"]["pageable_value"]), "ms/step", round(r["ms_per_step"],1),
              {k: round(v,1) for k,v in r["stage_ms"].items()}, "poa frac", round(r["roofline_poa"]["frac"],4), "lookup frac", round(r["roofline_lookup"]["frac"],3),
              "gaf", r.get("e2e_gaf",{}).get("value"), r.get("e2e_gaf",{}).get("gaf_mb_per_s"), "cpu", r.get("cpu_baseline",{}).get("value"), r.get("cpu_baseline",{}).get("t1_value"), "parity", r.get("parity_sample"))
PY
