#!/bin/bash
O=gpurun_out; tag=${1:-t}
( time timeout 1200 python -m pytest tests -m gpu -x -q -k "large_block or 216 or run_module or capacity" ) > $O/${tag}_pytest_big.log 2>&1; echo "pytest rc=$?" >> $O/${tag}_pytest_big.log
tail -12 $O/${tag}_pytest_big.log
timeout 400 python tools/bigpipe.py lifting > $O/${tag}_bigpipe.json 2> $O/${tag}_bigpipe.err; cat $O/${tag}_bigpipe.json; tail -3 $O/${tag}_bigpipe.err
