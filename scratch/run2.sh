./scratch/mb_fft > gpurun_out/mb_fft.txt 2>&1; cat gpurun_out/mb_fft.txt
for v in 3 2; do
B200_BR_KERNEL=$v timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_blind_rotate3' -s 1 -c 1 -f -o gpurun_out/v3_$v python scratch/prof_one.py 4736 > gpurun_out/v3_${v}_ncu.log 2>&1; tail -3 gpurun_out/v3_${v}_ncu.log
done
