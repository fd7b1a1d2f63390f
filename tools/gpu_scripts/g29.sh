mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_single.py tests/test_gpu_sharded.py -m gpu -q -x 2>&1 | tail -3
timeout 300 python bench.py --only c3 2>&1 | tail -1 | cut -c150-200
timeout 300 python tools/pipe_exp.py 100 2>&1 | tail -5
timeout 300 python bench.py --only c3s 2>&1 | tail -1 | cut -c190-290
