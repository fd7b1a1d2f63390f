mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_single.py -m gpu -q -x 2>&1 | tail -2
timeout 300 python bench.py --only c3 2>&1 | tail -1 | cut -c1-260
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_front --launch-skip 150 -c 1 -o gpurun_out/front_r1d -f python bench.py --only c3 > gpurun_out/ncu_front.log 2>&1
tail -1 gpurun_out/ncu_front.log | cut -c1-200
