run() { python bench.py --steps 300 --warmup 20 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['step_ms']['p50'], d['e2e']['ms_per_step'])"; }
for cap in 12 24 36 48 60; do echo "== wgcap=$cap"; PDL_WG_MAX_CTAS=$cap run; done
echo "== prio=2 wgcap=48"; PDL_CONVNET_PRIORITY=2 PDL_WG_MAX_CTAS=48 run
echo "== prio=2 wgcap=0"; PDL_CONVNET_PRIORITY=2 run
