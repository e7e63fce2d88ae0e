#!/bin/bash
# round-end check on one GPU: the full GPU suite, smoke(), the default bench line
O=gpurun_out; tag=${1:-final}
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${tag}_pytest_gpu.log
timeout 200 python __graft_entry__.py smoke >> $O/${tag}_pytest_gpu.log 2>&1; echo "smoke rc=$?" >> $O/${tag}_pytest_gpu.log
tail -12 $O/${tag}_pytest_gpu.log
timeout 900 python bench.py > $O/${tag}_bench_n1_lifting.json 2> $O/${tag}_bench_n1.err; echo "bench rc=$?"
head -c 300 $O/${tag}_bench_n1_lifting.json
