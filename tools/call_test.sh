#!/bin/bash
# developer GPU call: full GPU test suite, then the kernel A/B harness on the in-tree library
O=gpurun_out; tag=${1:-t}
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > $O/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${tag}_pytest.log
timeout 100 python tools/kbench.py 960 lifting 10 > $O/${tag}_kbench.json 2> $O/${tag}_kbench.err
tail -4 $O/${tag}_pytest.log; cat $O/${tag}_kbench.json
