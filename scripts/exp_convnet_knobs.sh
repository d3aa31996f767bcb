run1() { python bench.py --steps 300 --warmup 20 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['step_ms']['p50'], d['e2e']['ms_per_step'])"; }
python -m pytest tests/test_gpu_tc.py tests/test_gpu_tc_strided.py tests/test_gpu_guards.py tests/test_gpu_engine.py -m gpu -q 2>&1 | grep -E "^E  |passed|failed" | head -20
echo "== N=1 convnet"; run1
echo "== r50"; python bench.py --model resnet50 --steps 20 --warmup 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['e2e']['ms_per_step'])"
echo "== r18"; python bench.py --model resnet18 --steps 50 --warmup 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['e2e']['ms_per_step'])"
echo "== b4096"; python bench.py --batch 4096 --steps 30 --warmup 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'])"
python scripts/prof_timeline.py 100 gpurun_out/timeline_w1.txt > /dev/null 2>&1; cat gpurun_out/timeline_w1.txt
