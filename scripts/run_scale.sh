#!/bin/bash
# usage: scripts/run_scale.sh N   -- on an N-GPU box: multi-GPU tests, allreduce sweep, bench (both arms)
N=$1
mkdir -p gpurun_out
export PDL_KERNEL_TIMEOUT_S=20
timeout 400 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -4 > gpurun_out/multi_w$N.log; cat gpurun_out/multi_w$N.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench/allreduce_sweep.py --max-mb 256 > gpurun_out/sweep_w$N.log 2>&1; tail -14 gpurun_out/sweep_w$N.log
for impl in reference ours; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29577 bench.py --gpus $N --steps 300 --warmup 30 --impl $impl 2>&1 | tail -1 > gpurun_out/bench_${impl}_w$N.json; cut -c1-330 gpurun_out/bench_${impl}_w$N.json
done
for impl in reference ours; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29599 bench.py --gpus $N --model resnet50 --steps 20 --warmup 5 --impl $impl 2>&1 | tail -1 > gpurun_out/bench_r50_${impl}_w$N.json; cut -c1-330 gpurun_out/bench_r50_${impl}_w$N.json
done
grep -o '"exposed_comm_ms_per_step": [0-9.e-]*\|"compute_only_ms_per_step": [0-9.e-]*' gpurun_out/bench_*_w$N.json
