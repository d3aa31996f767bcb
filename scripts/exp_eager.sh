p() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], round(d['ms_per_step'],4), round(d['e2e']['ms_per_step'],4), d['config']['execution'][:30])" "$1"; }
python bench.py --model resnet18 --eager --steps 30 --warmup 5 2>/dev/null | tee gpurun_out/bench_r18_eager_w1.json | p r18_eager
python bench.py --model resnet18 --six-call --steps 50 --warmup 5 2>/dev/null | tee gpurun_out/bench_r18_sixcall_w1.json | p r18_sixcall
python bench.py --model resnet18 --impl reference --steps 50 --warmup 5 2>/dev/null | tee gpurun_out/bench_r18_ref_w1.json | p r18_ref
python bench.py --eager --steps 100 --warmup 10 2>/dev/null | tee gpurun_out/bench_eager_w1.json | p convnet_eager
