#!/bin/bash
# strong-scaling bench line on N GPUs of one box: tools/call_scale.sh <tag> <N>
O=gpurun_out; tag=$1; N=$2
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 --no-e2e > $O/${tag}_bench_n${N}_lifting.json 2> $O/${tag}_bench_n${N}.err; echo "rc=$?"
head -c 700 $O/${tag}_bench_n${N}_lifting.json; tail -3 $O/${tag}_bench_n${N}.err
