#!/bin/bash
# developer GPU call: source-level ncu capture of one kernel: call_ncu_one.sh <tag> <kernel regex> <bench args...>
O=gpurun_out; tag=$1; k=$2; shift 2
timeout 300 ncu --section SourceCounters --section WarpStateStats --section InstructionStats --clock-control none --import-source on -k regex:$k -c 1 -s 1 -o $O/${tag} python bench.py "$@" --steps 1 --warmup 1 --no-cpu --no-e2e > $O/${tag}_ncu.log 2>&1
tail -3 $O/${tag}_ncu.log
