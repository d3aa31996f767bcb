T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
run2() { $T --master-port $1 bench.py --gpus 2 --steps 100 --warmup 10 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['step_ms']['p50'], d['exposed_comm_ms_per_step'], d['cross_rank_ok'])"; }
run1() { python bench.py --steps 300 --warmup 20 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['step_ms']['p50'], d['e2e']['ms_per_step'])"; }
echo "== N=2 default"; run2 29511
echo "== N=2 dgrad2 cap 100"; PDL_CONVNET_DGRAD2_CTAS=100 run2 29512
echo "== N=2 wg2 main"; PDL_CONVNET_WG2_MAIN=1 run2 29513
echo "== N=2 wg2 main + wg cap 96"; PDL_CONVNET_WG2_MAIN=1 PDL_CONVNET_WG_CTAS=96 run2 29514
echo "== N=1 wg2 main"; PDL_CONVNET_WG2_MAIN=1 run1
