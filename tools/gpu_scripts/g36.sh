ASB_BATCH_VERBOSE=1 timeout 600 python bench.py --kind rot --quat-mode scalar_first --seqs 296 --frames 50 --steps 1 --warmup 1 --no-extras --cpu-seconds 1 2>&1 | tail -2 | cut -c1-600
ASB_BATCH_VERBOSE=1 timeout 600 python bench.py --kind ekf --markers 64 --p-visible 0.25 --seqs 592 --frames 200 --steps 1 --warmup 1 --no-extras --cpu-seconds 1 2>&1 | tail -2 | cut -c1-300
