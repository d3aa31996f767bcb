run1() { python bench.py --steps 400 --warmup 20 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],5), round(d['step_ms']['p50'],5), round(d['e2e']['ms_per_step'],5))"; }
for i in 1 2; do echo "== split"; run1; echo "== joined"; PDL_CONVNET_JOIN_C32=1 run1; done
