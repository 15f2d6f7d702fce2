# multi-GPU lines at N = 8 (gpurun --gpus 8): fit (image-parallel), decode8k (tile-row sharded), kodak (24 fits dealt to 8 ranks)
mkdir -p gpurun_out
n=8
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --no-cpu-baseline > gpurun_out/scale_fit_n$n.log 2>&1; tail -1 gpurun_out/scale_fit_n$n.log | cut -c1-300
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --workload decode8k --no-cpu-baseline > gpurun_out/scale_decode_n$n.log 2>&1; tail -1 gpurun_out/scale_decode_n$n.log | cut -c1-900
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $n --workload kodak --steps 300 > gpurun_out/scale_kodak_n$n.log 2>&1; tail -1 gpurun_out/scale_kodak_n$n.log | cut -c1-300
