mkdir -p gpurun_out
n=${1:-2}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --workload decode8k --no-cpu-baseline > gpurun_out/scale_decode_n$n.log 2>&1; tail -1 gpurun_out/scale_decode_n$n.log | cut -c1-1100
