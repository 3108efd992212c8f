#!/bin/bash
# A/B of the pair kernel's group width / blocks per lane on C4 (2 M reads) and C2; parity of the short-read tests first.
TAG=${1:-r02d}; O=gpurun_out; mkdir -p $O
for G in 8 4 2; do
  VGA_POA_G=$G timeout 600 python -m pytest tests -m gpu -q -x -k "c1 or hla or synthetic_short or edge or c2_full or cyclic or batch_ind or size_ind" > $O/${TAG}_pytest_G$G.log 2>&1; echo "pytest G=$G rc=$?"; tail -2 $O/${TAG}_pytest_G$G.log
done
for W in C4 C2; do
for G in 8 4 2; do for NB in 2 3; do
  R=""; [ $W = C4 ] && R="--reads 2000000"
  VGA_POA_G=$G VGA_POA_NBMAX=$NB timeout 300 python bench.py --workload $W $R --steps 3 --warmup 2 --no-e2e --no-gaf --no-cpu-baseline > $O/${TAG}_ab_${W}_G${G}_NB${NB}.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("$O/${TAG}_ab_${W}_G${G}_NB${NB}.json"))
print("$W G=$G NB=$NB", "ms_poa %.1f" % d["stage_ms"]["ms_poa"], "ms_total %.1f" % d["stage_ms"]["ms_total"], "frac %.4f" % d["roofline_poa"]["frac"])
PY
done; done; done
