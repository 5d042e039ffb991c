# Scaling runs of the round on one 8-GPU box (one process per GPU under torchrun, no data-path
# collective): the full bench line (MQ + HT decode, configs[3] tile-sharded, encode legs frame-sharded)
# at N = 2, 4, 8.  Lines go to gpurun_out/scale_n*.json.  The N = 1 line is the profile pass's.
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29600+n)) bench.py --gpus $n > gpurun_out/scale_n$n.json 2>> gpurun_out/scale_err.log
done
tail -5 gpurun_out/scale_err.log
for f in gpurun_out/scale_n*.json; do echo $f; cut -c1-300 $f; done
