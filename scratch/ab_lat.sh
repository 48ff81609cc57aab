for m in ${LAT_MODES:-1 3}; do B200_LAT=$m timeout 300 python scratch/ab_lat.py 80 2>&1 | tail -1; done
for m in ${LAT_MODES:-1 3}; do B200_LAT=$m timeout 300 python scratch/ab_lat.py 128 2>&1 | tail -1; done
