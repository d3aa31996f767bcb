#!/usr/bin/env bash
# Experiment driver (round 2): tc_conv as CTA pairs with a multicast weight tile (PDL_TC_CLUSTER) -- numerics with pairs forced on,
# then ResNet-50 and ConvNet b=4096 with and without.  Results: profiles/tc_conv_cta_pairs.md.
echo "== forced pairs on the test shapes"; PDL_TC_CLUSTER=2 python -m pytest tests/test_gpu_tc.py tests/test_gpu_tc_strided.py tests/test_gpu_ops.py -m gpu -x -q -k "not emulated" 2>&1 | tail -3
echo "== default"; python -m pytest tests/test_gpu_tc.py tests/test_gpu_tc_strided.py tests/test_gpu_engine.py -m gpu -x -q 2>&1 | tail -2
for c in 0 1; do echo "== resnet50 cluster=$c"; PDL_TC_CLUSTER=$c python bench.py --model resnet50 --steps 20 --warmup 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['e2e']['ms_per_step'])"; done
for c in 0 1; do echo "== convnet b4096 cluster=$c"; PDL_TC_CLUSTER=$c python bench.py --batch 4096 --steps 30 --warmup 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'])"; done
