#!/bin/bash
# One GPU-box visit: parity tests, bench line (both arms), ncu launch list, DRAM traffic and full captures of the two
# hot kernels.   usage (under gpurun): bash tools/gpu_round.sh <tag> [skip-tests]
# Everything lands in gpurun_out/<tag>_*; copy what should be judged into profiles/.
set -u
TAG=${1:-run}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
nproc > $OUT/${TAG}_host.txt; lscpu | grep -E "Model name|^CPU\(s\)" >> $OUT/${TAG}_host.txt
if [ "${2:-}" != "skip-tests" ]; then
  timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1
  echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
  tail -5 $OUT/${TAG}_pytest.log
fi
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench.err
timeout 900 python bench.py --steps 5 --warmup 3 > $OUT/${TAG}_bench.json 2>> $OUT/${TAG}_bench.err
echo "bench exit $?"; head -c 700 $OUT/${TAG}_bench.json; echo
# launch list of a short bench run (shares of the step); -s skips key creation and warm-up launches
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-p128 > $OUT/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv \
    --log-file $OUT/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-p128 > $OUT/${TAG}_ncu1.log 2>&1
# DRAM traffic of the two hot kernels at the bench's own launch size (65,536 gates)
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,launch__grid_size --clock-control none \
    -k regex:'k_blind_rotate|k_keyswitch' -s 2 -c 2 -f -o $OUT/${TAG}_traffic python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-p128 > $OUT/${TAG}_ncu3.log 2>&1
# full capture of one launch of each hot kernel (8,880 gates = 10 whole waves of the six-ciphertext blind rotation, one launch)
timeout 300 python scratch/prof_one.py 8880 > $OUT/${TAG}_prof_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_blind_rotate|k_keyswitch' -s 2 -c 2 \
    -f -o $OUT/${TAG}_prof python scratch/prof_one.py 8880 > $OUT/${TAG}_ncu2.log 2>&1
tail -3 $OUT/${TAG}_ncu2.log
ls -la $OUT | tail -12
# small batches, secondary workloads (BASELINE configs[2], [3]) and one full capture of the cluster latency kernel
timeout 300 python scratch/small_batches.py > $OUT/${TAG}_small_batches.txt 2>&1
timeout 1200 python tools/workloads.py all > $OUT/${TAG}_workloads.jsonl 2> $OUT/${TAG}_workloads.err
timeout 300 python scratch/prof_one.py 1 > $OUT/${TAG}_lat2_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_blind_rotate_lat2' -s 1 -c 1 \
    -f -o $OUT/${TAG}_lat2 python scratch/prof_one.py 1 > $OUT/${TAG}_ncu4.log 2>&1
tail -2 $OUT/${TAG}_ncu4.log
