run1() { python bench.py --steps 300 --warmup 20 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['step_ms']['p50'], d['e2e']['ms_per_step'])"; }
echo "== default"; run1
echo "== wg2 prio"; PDL_CONVNET_WG2_PRIO=1 run1
echo "== wg2 prio + dgrad2 cap 100"; PDL_CONVNET_WG2_PRIO=1 PDL_CONVNET_DGRAD2_CTAS=100 run1
echo "== wg2 prio + wg ctas 32"; PDL_CONVNET_WG2_PRIO=1 PDL_CONVNET_WG_CTAS=33 run1
python -m pytest tests/test_gpu_guards.py tests/test_gpu_tc.py -m gpu -x -q 2>&1 | tail -3
