mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:^k_batch$" -c 1 -o gpurun_out/batch_r01e -f python bench.py --seqs 444 --frames 100 --steps 1 --warmup 0 --no-extras --cpu-seconds 1 > /dev/null 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_c2_r01e.csv python bench.py --seqs 444 --frames 100 --steps 1 --warmup 1 --no-extras --cpu-seconds 1 > /dev/null 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 400 -c 45 --csv --log-file gpurun_out/launches_c3_r01e.csv python bench.py --only c3 > /dev/null 2>&1
ls -la gpurun_out/*r01e*
