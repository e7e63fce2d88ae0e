#!/bin/bash
O=gpurun_out; tag=${1:-t}
( time timeout 1200 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -k "large_blocks_full or mid_size" --durations=8 ) > $O/${tag}_pytest_sub.log 2>&1; echo "pytest rc=$?" >> $O/${tag}_pytest_sub.log
tail -22 $O/${tag}_pytest_sub.log
