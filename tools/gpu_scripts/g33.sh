# round-1 final evidence: full tests, smoke, bench line, launch lists, full captures
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
timeout 900 python bench.py > gpurun_out/bench_r01c.json 2> gpurun_out/bench_r01c.err; tail -c 400 gpurun_out/bench_r01c.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 400 -c 45 --csv --log-file gpurun_out/launches_c3_r01c.csv python bench.py --only c3 > /dev/null 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_c2_r01c.csv python bench.py --seqs 444 --frames 100 --steps 1 --warmup 1 --no-extras --cpu-seconds 1 > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:^k_batch$" -c 1 -o gpurun_out/batch_r01c -f python bench.py --seqs 444 --frames 100 --steps 1 --warmup 0 --no-extras --cpu-seconds 1 > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_sweep --launch-skip 150 -c 1 -o gpurun_out/sweep_r01c -f python bench.py --only c3 > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_f2 --launch-skip 150 -c 1 -o gpurun_out/f2_r01c -f python bench.py --only c3 > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_f1 --launch-skip 150 -c 1 -o gpurun_out/f1_r01c -f python bench.py --only c3 > /dev/null 2>&1
ls -la gpurun_out/*r01c*
