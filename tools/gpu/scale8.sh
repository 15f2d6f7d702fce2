# multi-GPU scaling lines: fit (image-parallel) and decode8k (tile-row sharded) at N = 8 and 4
mkdir -p gpurun_out
for n in 8 4; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --no-cpu-baseline > gpurun_out/scale_fit_n$n.log 2>&1; tail -1 gpurun_out/scale_fit_n$n.log | cut -c1-600
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --workload decode8k --no-cpu-baseline > gpurun_out/scale_decode_n$n.log 2>&1; tail -1 gpurun_out/scale_decode_n$n.log | cut -c1-900
done
