mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_single.py -m gpu -q -x 2>&1 | tail -15
timeout 600 python bench.py --only c3 2>&1 | tail -1 | cut -c1-700
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 400 -c 40 --csv --log-file gpurun_out/launches_c3_r1c.csv python bench.py --only c3 > gpurun_out/ncu_c3d.log 2>&1
tail -2 gpurun_out/ncu_c3d.log | cut -c1-300
