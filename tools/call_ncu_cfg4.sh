#!/bin/bash
# developer GPU call: source-level ncu capture of the config-4 kernels on a 128-plane slab
O=gpurun_out; tag=${1:-c4}
timeout 300 python bench.py --config 4 --nz 128 --steps 3 --warmup 1 --no-cpu --no-e2e > $O/${tag}_bench.json 2> $O/${tag}_bench.err
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_encode_fused -c 1 -s 1 -o $O/${tag}_enc python bench.py --config 4 --nz 128 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/${tag}_ncu_enc.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_decode_fused -c 1 -s 1 -o $O/${tag}_dec python bench.py --config 4 --nz 128 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/${tag}_ncu_dec.log 2>&1
cat $O/${tag}_bench.json | head -c 1500
