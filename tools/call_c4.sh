#!/bin/bash
O=gpurun_out; tag=${1:-t}
timeout 300 python bench.py --config 4 --steps 5 --warmup 2 --no-cpu --no-e2e > $O/${tag}_c4.json 2> $O/${tag}_c4.err; echo "c4 rc=$?"
python -c "
import json,sys
d=json.load(open('$O/${tag}_c4.json')); print(d['value'], d['compress_ms'], d['decompress_ms'], d['compressed_length'], d['snr_db'], {k:(round(v['ms'],3), round(v['frac'],3)) for k,v in d['roofline']['kernels'].items()})
"; tail -3 $O/${tag}_c4.err
