timeout 300 python tools/sweep_exp.py 0 2 1 3 2>&1 | grep "config 1"
