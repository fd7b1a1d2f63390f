mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_single.py -m gpu -q -x 2>&1 | tail -4
python tools/sweep_exp.py 0 2>&1 | tail -3
timeout 600 python bench.py --only c3 2>&1 | tail -1 | cut -c1-300
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_front --launch-skip 150 -c 1 -o gpurun_out/front_r1c -f python bench.py --only c3 > gpurun_out/ncu_front.log 2>&1
tail -2 gpurun_out/ncu_front.log | cut -c1-200
