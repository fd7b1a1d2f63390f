for pad in 0 1 2 3 4 6 8 13; do echo "pad $pad x16 doubles"; ASB_LD_PAD=$pad timeout 300 python tools/sweep_exp.py 2 3 2>&1 | grep "^config 1"; done
