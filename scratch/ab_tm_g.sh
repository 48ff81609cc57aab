#!/bin/bash
# A/B of the tm kernel at 4 / 5 / 6 ciphertexts per CTA
for g in 6 5 4; do echo "== B200_TM_G=$g"; B200_TM_G=$g timeout 300 python scratch/ab_tm.py 80 8880 65536 2>&1 | grep -v "^P80 v2"; done
timeout 300 python scratch/ab_tm.py 128 5920
