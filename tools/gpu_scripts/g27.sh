mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -q -x 2>&1 | tail -3
timeout 600 python bench.py --only c3s 2>&1 | tail -1 | cut -c190-290
ASB_TAIL=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --only c3s 2>&1 | tail -1 | cut -c190-290
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --only c3s 2>&1 | tail -1 | cut -c190-290
