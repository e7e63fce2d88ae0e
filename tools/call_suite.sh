#!/bin/bash
# developer GPU call: the full GPU test suite + smoke
O=gpurun_out; tag=${1:-t}
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > $O/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${tag}_pytest.log
timeout 200 python __graft_entry__.py smoke >> $O/${tag}_pytest.log 2>&1; echo "smoke rc=$?" >> $O/${tag}_pytest.log
tail -12 $O/${tag}_pytest.log
