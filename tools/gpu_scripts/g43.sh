# final check of the committed tree: all GPU tests, smoke, the default bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
timeout 900 python bench.py > gpurun_out/bench_r01e.json 2> gpurun_out/bench_r01e.err; tail -c 300 gpurun_out/bench_r01e.json; tail -3 gpurun_out/bench_r01e.err
