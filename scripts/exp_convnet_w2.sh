T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
run2() { $T --master-port $1 bench.py --gpus 2 --steps 200 --warmup 10 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],5), round(d['step_ms']['p50'],5), round(d['exposed_comm_ms_per_step'],5), d['cross_rank_ok'])"; }
run1() { python bench.py --steps 400 --warmup 20 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],5), round(d['step_ms']['p50'],5), round(d['e2e']['ms_per_step'],5))"; }
python -m pytest tests/test_gpu_multi.py -m gpu -q -k "graph or sixcall" 2>&1 | tail -2
echo "== N=1"; run1
echo "== N=2 split c3/c2"; run2 29511
echo "== N=2 joined"; PDL_CONVNET_JOIN_C32=1 run2 29512
