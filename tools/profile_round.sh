set -x
python bench.py --steps 2 --warmup 3 > gpurun_out/p_bench.json 2> gpurun_out/p_bench.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/p_launches_j2k97.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/p_ncu1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/p_launches_ht.csv python bench.py --workload ht --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/p_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"idwt_strip|t1_mq_decode" --launch-skip 6 -c 6 -o gpurun_out/p_full_j2k97 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/p_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"idwt_strip|t1_ht_decode" --launch-skip 6 -c 6 -o gpurun_out/p_full_ht python bench.py --workload ht --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/p_ncu4.log 2>&1
python tools/bench_extra.py ht_encode --steps 2 --no-cpu > gpurun_out/p_ht_encode.json 2>> gpurun_out/p_bench.err
ncu --set full --clock-control none --import-source on -k regex:"fdwt_strip|t1_ht_encode|pack_" --launch-skip 7 -c 7 -o gpurun_out/p_full_ht_encode python tools/bench_extra.py ht_encode --steps 1 --no-cpu > gpurun_out/p_ncu5.log 2>&1
ls -la gpurun_out
