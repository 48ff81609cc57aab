for m in 0 6 1 2 3 4 5 7; do B200_KS_MODE=$m timeout 300 python scratch/ab_ks.py 65536 2>&1 | tail -1; done
