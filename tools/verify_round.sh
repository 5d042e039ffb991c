# Round-end verification on one B200: GPU tests, smoke(), the two bench.py workloads, the encode
# extras and a fresh --set full capture of the two MQ encode kernels.  Run under gpurun:
#   gpurun --timeout 1500 -- 'bash tools/verify_round.sh'
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/v_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/v_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/v_smoke.log 2>&1
timeout 400 python bench.py > gpurun_out/v_bench_j2k97.json 2> gpurun_out/v_bench.err
timeout 400 python bench.py --workload ht > gpurun_out/v_bench_ht.json 2>> gpurun_out/v_bench.err
timeout 300 python tools/bench_extra.py ht_encode > gpurun_out/v_ht_encode.json 2>> gpurun_out/v_bench.err
timeout 300 python tools/bench_extra.py mq_encode --frames 16 --inflight 8 > gpurun_out/v_mq_encode.json 2>> gpurun_out/v_bench.err
timeout 300 python tools/bench_extra.py ht_encode --samples16 --no-cpu > gpurun_out/v_ht_encode_s16.json 2>> gpurun_out/v_bench.err
timeout 300 python tools/bench_extra.py mq_encode --frames 16 --inflight 8 --samples16 --no-cpu > gpurun_out/v_mq_encode_s16.json 2>> gpurun_out/v_bench.err
timeout 400 python tools/bench_extra.py plugin_batch --frames 48 --steps 1 > gpurun_out/v_plugin_batch.json 2>> gpurun_out/v_bench.err
timeout 400 python tools/bench_extra.py plugin_batch_decode --frames 48 --steps 1 --inflight 8 > gpurun_out/v_plugin_batch_decode.json 2>> gpurun_out/v_bench.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/v_launches_mq_encode.csv python tools/bench_extra.py mq_encode --frames 2 --inflight 1 --steps 1 --no-cpu > gpurun_out/v_ncu1.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"t1_mq_model|t1_mq_code|t1_mq_order" --launch-skip 3 -c 3 -o gpurun_out/v_full_mq_encode python tools/bench_extra.py mq_encode --frames 2 --inflight 1 --steps 1 --no-cpu > gpurun_out/v_ncu2.log 2>&1
tail -3 gpurun_out/v_pytest.log; cat gpurun_out/v_smoke.log | tail -1; ls -la gpurun_out | tail -15
