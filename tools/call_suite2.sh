#!/bin/bash
O=gpurun_out; tag=${1:-final}
( time timeout 1500 python -m pytest tests -m gpu -x -q --durations=6 ) > $O/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${tag}_pytest_gpu.log
timeout 200 python __graft_entry__.py smoke >> $O/${tag}_pytest_gpu.log 2>&1; echo "smoke rc=$?" >> $O/${tag}_pytest_gpu.log
tail -22 $O/${tag}_pytest_gpu.log
