#!/bin/bash
# round-end measurement call on one GPU: config-3 bench line, launch list, ncu --set full raw metrics of the three hot kernels,
# then the bench lines of configs 1, 2 and 5
O=gpurun_out; tag=${1:-r02_v5}
timeout 900 python bench.py --steps 10 --warmup 3 > $O/${tag}_bench_n1_lifting.json 2> $O/${tag}_bench_n1.err; echo "bench rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${tag}_launches_lifting.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > $O/${tag}_ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:'k_sumsq_rows|k_encode32ws|k_decode32ws' -c 3 -s 6 -o /tmp/${tag}_full python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > $O/${tag}_ncu_full.log 2>&1
ncu -i /tmp/${tag}_full.ncu-rep --page raw --csv > $O/${tag}_ncu_raw_lifting.csv 2>> $O/${tag}_ncu_full.log
for c in 1 2 5; do timeout 900 python bench.py --config $c --steps 10 --warmup 3 > $O/${tag}_config${c}_n1.json 2> $O/${tag}_config${c}_n1.err; echo "config $c rc=$?"; done
head -c 400 $O/${tag}_bench_n1_lifting.json; echo; ls -la $O
