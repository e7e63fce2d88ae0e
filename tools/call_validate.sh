O=gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/r02v3_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02v3_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/r02v3_bench_n1.json 2> $O/r02v3_bench_n1.err; echo "bench rc=$?" >> $O/r02v3_bench_n1.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02v3_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > $O/r02v3_ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_encode32ws -c 1 -s 2 -o $O/r02v3_enc python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > $O/r02v3_ncu_enc.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_decode32ws -c 1 -s 2 -o $O/r02v3_dec python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > $O/r02v3_ncu_dec.log 2>&1
tail -3 $O/r02v3_pytest.log; cat $O/r02v3_bench_n1.json | head -c 600
