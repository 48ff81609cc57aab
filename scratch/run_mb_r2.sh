#!/bin/bash
# round-2 pipe probes: DMMA, TMEM exchange, lane map
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2_mb_smi.txt
timeout 120 scratch/mb_r2 map  > gpurun_out/r2_mb_map.txt 2>&1;  echo "map exit $?"
timeout 300 scratch/mb_r2 dmma > gpurun_out/r2_mb_dmma.txt 2>&1; echo "dmma exit $?"
timeout 300 scratch/mb_r2 tmem > gpurun_out/r2_mb_tmem.txt 2>&1; echo "tmem exit $?"
cat gpurun_out/r2_mb_dmma.txt gpurun_out/r2_mb_tmem.txt
head -50 gpurun_out/r2_mb_map.txt
