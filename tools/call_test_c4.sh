#!/bin/bash
# developer GPU call: config-4 bench (short), then the full GPU test suite
O=gpurun_out; tag=${1:-t}
timeout 300 python bench.py --config 4 --steps 5 --warmup 2 --no-cpu --no-e2e > $O/${tag}_c4.json 2> $O/${tag}_c4.err; echo "c4 rc=$?"
CVX_GPU_NO_SPARSE=1 timeout 300 python bench.py --config 4 --steps 5 --warmup 2 --no-cpu --no-e2e > $O/${tag}_c4_old.json 2> $O/${tag}_c4_old.err; echo "c4 old rc=$?"
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > $O/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${tag}_pytest.log
tail -15 $O/${tag}_pytest.log
python - <<'PY'
import json,sys
for f in ("gpurun_out/%s_c4.json" % sys.argv[1] if len(sys.argv)>1 else "", ):
    pass
PY
for f in $O/${tag}_c4.json $O/${tag}_c4_old.json; do python -c "
import json,sys
d=json.load(open('$f')); print('$f', d['value'], d['compress_ms'], d['decompress_ms'], d['compressed_length'], d['snr_db'], {k:(round(v['ms'],3), round(v['frac'],3)) for k,v in d['roofline']['kernels'].items()})
"; done; tail -3 $O/${tag}_c4.err
