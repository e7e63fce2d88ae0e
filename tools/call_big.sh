#!/bin/bash
O=gpurun_out; tag=${1:-t}
( time timeout 900 python -m pytest tests -m gpu -x -q -k "large_block or 216 or run_module" ) > $O/${tag}_pytest_big.log 2>&1; echo "pytest rc=$?" >> $O/${tag}_pytest_big.log
tail -6 $O/${tag}_pytest_big.log
timeout 600 python bench.py --config 2 --steps 10 --warmup 3 > $O/${tag}_config2_n1.json 2> $O/${tag}_config2_n1.err; echo "c2 rc=$?"
python -c "
import json; e=json.load(open('$O/${tag}_config2_n1.json')); print({k:(round(v['mcells_s']), round(v['frac'],3), v['max_roundtrip_error']) for k,v in e['sweep'].items()})"
