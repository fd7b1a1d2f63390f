mkdir -p gpurun_out
N=${1:-2}
timeout 600 python bench.py --only c3s 2>&1 | tail -1 | cut -c1-400 | tee gpurun_out/g21_c3s_w1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --only c3s 2>&1 | tail -2 | cut -c1-400 | tee gpurun_out/g21_c3s_w$N.json
