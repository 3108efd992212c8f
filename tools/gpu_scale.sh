#!/bin/bash
# the default bench under torchrun on N GPUs of one box (what the driver's scaling run does).  usage: tools/gpu_scale.sh <tag> <N>   (gpurun --gpus N)
O=gpurun_out; mkdir -p $O; TAG=${1:-rXX}; N=${2:-8}
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 --cpu-seconds 4 > $O/${TAG}_bench_default_n$N.json 2> $O/${TAG}_bench_default_n$N.err; echo "bench n=$N rc=$?"; tail -3 $O/${TAG}_bench_default_n$N.err
python - <<PY
import json
d=json.load(open("$O/${TAG}_bench_default_n$N.json"))
for name, r in (("C4", d), ("C5", d.get("long_reads"))):
    print(name, "n_gpus", d["n_gpus"], "value", round(r["value"]), "e2e", round(r["e2e"]["value"]), "ms/step", round(r["ms_per_step"],1), "bcast_s", round(r["setup"]["index_broadcast_s"],2), "parity", r.get("parity_sample"), "sharded", r.get("e2e_sharded"))
PY
