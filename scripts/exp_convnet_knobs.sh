#!/usr/bin/env bash
# Experiment driver (round 2): the ConvNet schedule knobs of ops/convnet_engine.py one by one on one B200; each line prints
# ms/step (mean, p50) and end-to-end ms/step.  Results are quoted in DESIGN.md 2c.
run1() { python bench.py --steps 400 --warmup 20 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],5), round(d['step_ms']['p50'],5), round(d['e2e']['ms_per_step'],5))"; }
echo "== default"; run1
echo "== fuse pool2"; PDL_CONVNET_FUSE_POOL2=1 run1
echo "== fuse fc"; PDL_CONVNET_FUSE_FC=1 run1
echo "== wg 40"; PDL_CONVNET_WG_CTAS=40 run1
echo "== wg 64"; PDL_CONVNET_WG_CTAS=64 run1
echo "== dgrad2 cap 0"; PDL_CONVNET_DGRAD2_CTAS=0 run1
echo "== no pdl"; PDL_NO_PDL=1 run1
