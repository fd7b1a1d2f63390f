set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_single.py -m gpu -q -x 2>&1 | tail -15
timeout 600 python bench.py --only c3 2>&1 | tail -3
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 700 -c 60 --csv --log-file gpurun_out/launches_c3_full.csv python bench.py --only c3 > gpurun_out/ncu_c3b.log 2>&1
