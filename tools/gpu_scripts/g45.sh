mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -q -x 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_r01e_n$N.json 2> gpurun_out/bench_r01e_n$N.err; tail -c 900 gpurun_out/bench_r01e_n$N.json
