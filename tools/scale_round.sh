# Scaling runs of the round on one 8-GPU box (one process per GPU under torchrun, no data-path
# collective): the headline workload at N=8, HTJ2K 4K decode at N=2,4,8, the 8K tiled decode and
# the DCI-4K batch encode at N=8.  Lines go to gpurun_out/scale_*.json.
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29600+n)) bench.py --workload ht --gpus $n --steps 10 --warmup 3 > gpurun_out/scale_ht_n$n.json 2>> gpurun_out/scale_err.log
done
$TR --nproc-per-node 8 --master-port 29620 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/scale_j2k97_n8.json 2>> gpurun_out/scale_err.log
$TR --nproc-per-node 8 --master-port 29621 tools/bench_extra.py tiled --steps 3 > gpurun_out/scale_tiled_n8.json 2>> gpurun_out/scale_err.log
$TR --nproc-per-node 8 --master-port 29622 tools/bench_extra.py mq_encode --steps 2 --frames 32 > gpurun_out/scale_mq_encode_n8.json 2>> gpurun_out/scale_err.log
tail -5 gpurun_out/scale_err.log
for f in gpurun_out/scale_*.json; do echo $f; cut -c1-400 $f; done
