for m in 0 2 1 3; do B200_LAT=$m timeout 300 python scratch/ab_lat.py 80 2>&1 | tail -2; done
for m in 2 1; do B200_LAT=$m timeout 300 python scratch/ab_lat.py 128 2>&1 | tail -2; done
