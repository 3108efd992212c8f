#!/bin/bash
# 2-GPU visit: multi-GPU parity through the C ABI (python threads + the C++ harness), bench under torchrun with vga_index_broadcast
O=gpurun_out; mkdir -p $O; TAG=${1:-r02q}; N=${2:-2}
nvidia-smi --query-gpu=name --format=csv,noheader | head -8
timeout 600 python -m pytest tests -m gpu -q -x -k "n_gpus" > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
python - <<'PY'
import sys; sys.path.insert(0, ".")
import vgaligner_b200 as vg
ix = vg.Index.build("tests/golden/graphs/19-MICB-4227-smooth.gfa", 11)
ix.store_in_file("/tmp/micb.vgidx")
r = vg.simulate_reads(ix, 20000, 150, 0.01, 0.0005, 0.0005, 0.1, 5)
import ctypes as C
vg.host_lib().vgah_reads_write_fasta(r._h, b"/tmp/micb_reads.fa")
PY
./rs-vgaligner_b200/vga_map map -i /tmp/micb.vgidx -f /tmp/micb_reads.fa -o /tmp/out1 --also-align 2>&1 | tail -1
./rs-vgaligner_b200/vga_map map -i /tmp/micb.vgidx -f /tmp/micb_reads.fa -o /tmp/outN --also-align --gpus $N 2>&1 | tail -1
cmp /tmp/out1-chains.gaf /tmp/outN-chains.gaf && cmp /tmp/out1-alignments.gaf /tmp/outN-alignments.gaf && echo "vga_map --gpus $N: GAF files identical to 1 GPU ($(wc -l < /tmp/outN-alignments.gaf) alignment lines)"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 2 --warmup 2 --cpu-seconds 4 > $O/${TAG}_bench_n$N.json 2> $O/${TAG}_bench_n$N.err; echo "bench n=$N rc=$?"; tail -3 $O/${TAG}_bench_n$N.err
python - <<PY
import json
d=json.load(open("$O/${TAG}_bench_n$N.json"))
for name, r in (("C4", d), ("C5", d.get("long_reads"))):
    print(name, "n_gpus", d["n_gpus"], "value", round(r["value"]), "e2e", round(r["e2e"]["value"]), "ms/step", round(r["ms_per_step"],1), "bcast_s", r["setup"]["index_broadcast_s"], "parity", r.get("parity_sample"))
PY
