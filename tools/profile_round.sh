# One-GPU profile pass of a round: bench lines, launch lists, full captures of the dominant kernels.
# Run on the GPU box through gpurun; everything lands in gpurun_out/ (copy what is to be kept to profiles/).
set -x
python bench.py > gpurun_out/p_bench.json 2> gpurun_out/p_bench.err || exit 1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/p_bench_ref.json 2>> gpurun_out/p_bench.err
Q="--no-plugin --no-cpu-baseline --no-encode --no-tiled"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/p_launches_bench.csv python bench.py --steps 2 --warmup 3 $Q > gpurun_out/p_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"idwt_strip|t1_mq_decode" --launch-skip 6 -c 6 -o gpurun_out/p_full_mq python bench.py --steps 1 --warmup 1 --no-ht $Q > gpurun_out/p_ncu2.log 2>&1
GRKB_T1_MQ_GROUP=16 ncu --set full --clock-control none --import-source on -k regex:"t1_mq_decode" --launch-skip 2 -c 1 -o gpurun_out/p_full_mq_pairs python bench.py --steps 1 --warmup 1 --no-ht $Q > gpurun_out/p_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"idwt_strip|t1_ht_decode" --launch-skip 6 -c 6 -o gpurun_out/p_full_ht python bench.py --workload ht --steps 1 --warmup 1 $Q > gpurun_out/p_ncu4.log 2>&1
python tools/bench_extra.py compress --steps 12 > gpurun_out/p_compress_dci.json 2>> gpurun_out/p_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/p_launches_compress.csv python tools/bench_extra.py compress --steps 1 --no-cpu > gpurun_out/p_ncu5.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"pcrd_step|pcrd_select" --launch-skip 70 -c 2 -o gpurun_out/p_full_pcrd python tools/bench_extra.py compress --steps 1 --no-cpu > gpurun_out/p_ncu6.log 2>&1
ls -la gpurun_out
