run1() { python bench.py --steps 300 --warmup 20 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['step_ms']['p50'], d['e2e']['ms_per_step'])"; }
python -m pytest tests/test_gpu_guards.py tests/test_gpu_engine.py tests/test_gpu_async_step.py tests/test_gpu_tc.py tests/test_gpu_multi.py tests/test_gpu_multi_failure.py tests/test_gpu_examples.py -m gpu -q 2>&1 | grep -E "^E  |passed|failed" | head -20
echo "== N=1 default"; run1
echo "== N=1 no fuse c2"; PDL_CONVNET_FUSE_C2=0 run1
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
run2() { $T --master-port $1 bench.py --gpus 2 --steps 100 --warmup 10 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['step_ms']['p50'], d['exposed_comm_ms_per_step'], d['cross_rank_ok'])"; }
echo "== N=2 default"; run2 29511
echo "== N=2 no fuse c2"; PDL_CONVNET_FUSE_C2=0 run2 29512
$T --master-port 29502 scripts/prof_timeline.py 100 gpurun_out/timeline_w2.txt > /dev/null 2> gpurun_out/timeline_w2.err; head -24 gpurun_out/timeline_w2.txt
