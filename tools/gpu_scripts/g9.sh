mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_sweep --launch-skip 150 -c 1 -o gpurun_out/sweep_r1c -f python bench.py --only c3 > gpurun_out/ncu_sweep.log 2>&1
tail -2 gpurun_out/ncu_sweep.log | cut -c1-200
