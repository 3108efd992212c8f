#!/bin/bash
O=gpurun_out; mkdir -p $O; TAG=${1:-r02m}
B="--no-cpu-baseline --no-e2e --no-gaf --no-index-build --no-sharded"
env VGA_HOST_TIMELINE=1 timeout 300 python bench.py --steps 1 --warmup 1 --workload C5 --reads 50000 $B > $O/${TAG}_n1.json 2> $O/${TAG}_n1.err; echo "n1 rc=$?"
grep "fused\|call done\|pool ready" $O/${TAG}_n1.err | tail -12
env VGA_HOST_TIMELINE=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29523 --log-dir $O/${TAG}_logs --redirects 3 bench.py --gpus 2 --steps 1 --warmup 1 --workload C5 $B > /dev/null 2>&1; echo "n2 rc=$?"
for r in 0 1; do f=$(find $O/${TAG}_logs -name stderr.log -path "*/$r/*" | head -1); echo "rank $r"; grep "fused\|call done\|pool ready" $f | tail -12; done
nvidia-smi --query-gpu=index,clocks.mem,clocks.sm,memory.used,memory.total --format=csv
