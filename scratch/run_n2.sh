timeout 600 python -m pytest tests/test_gpu_workloads.py -q -m gpu -k "multi_device" 2>&1 | tail -2
echo "NCCL_DEBUG in the box environment: [${NCCL_DEBUG:-unset}]"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err
echo "exit $?"; tail -c 2500 gpurun_out/r2_bench_n2.json; grep -iE "nranks|NVLS|error" gpurun_out/r2_bench_n2.err | head -8
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 | head -c 400
