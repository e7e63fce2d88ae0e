#!/bin/bash
O=gpurun_out; tag=${1:-t}
( time timeout 900 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -k "mid_size" ) > $O/${tag}_pytest_mid.log 2>&1; echo "pytest rc=$?" >> $O/${tag}_pytest_mid.log
tail -8 $O/${tag}_pytest_mid.log
timeout 900 python bench.py --config 4 --steps 10 --warmup 3 > $O/${tag}_config4_n1.json 2> $O/${tag}_config4_n1.err; echo "c4 rc=$?"
head -c 600 $O/${tag}_config4_n1.json; tail -3 $O/${tag}_config4_n1.err
