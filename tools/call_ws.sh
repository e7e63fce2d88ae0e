#!/bin/bash
# developer GPU call: the tests that exercise the warp-specialised 32^3 kernels, smoke, and the kernel A/B harness
O=gpurun_out; tag=${1:-t}
( time timeout 600 python -m pytest tests -m gpu -x -q -k "warp_specialised or lifting or golden_config1 or nan_and_inf or out_of_range or two_cuda_streams or torch_ops" ) > $O/${tag}_pytest_ws.log 2>&1; echo "pytest rc=$?" >> $O/${tag}_pytest_ws.log
timeout 200 python __graft_entry__.py smoke >> $O/${tag}_pytest_ws.log 2>&1; echo "smoke rc=$?" >> $O/${tag}_pytest_ws.log
tail -8 $O/${tag}_pytest_ws.log
timeout 100 python tools/kbench.py 960 lifting 10 > $O/${tag}_kbench.json 2> $O/${tag}_kbench.err; cat $O/${tag}_kbench.json
