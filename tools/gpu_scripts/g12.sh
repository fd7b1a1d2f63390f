mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:^k_batch$" -c 1 -o gpurun_out/batch_r1b -f python bench.py --seqs 1184 --frames 40 --steps 1 --warmup 0 --no-extras --cpu-seconds 1 > gpurun_out/ncu_c2b.log 2>&1
tail -1 gpurun_out/ncu_c2b.log | cut -c1-200
