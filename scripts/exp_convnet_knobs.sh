run() { python bench.py --steps 300 --warmup 20 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['step_ms']['p50'], d['e2e']['ms_per_step'])"; }
echo "== default"; run
echo "== dgrad2 cap 100"; PDL_CONVNET_DGRAD2_CTAS=100 run
echo "== wg 45"; PDL_CONVNET_WG_CTAS=45 run
echo "== wg 51"; PDL_CONVNET_WG_CTAS=51 run
python -m pytest tests/test_gpu_tc.py tests/test_gpu_engine.py tests/test_gpu_async_step.py -m gpu -x -q 2>&1 | tail -2
