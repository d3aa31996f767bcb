#!/usr/bin/env bash
# The round's table: both arms back to back on one 8 x B200 box.  Every JSON line lands in gpurun_out/scale/.
mkdir -p gpurun_out/scale
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
port=29700
run() {  # name nproc args...
  local name=$1 n=$2; shift 2
  port=$((port + 1))
  if [ "$n" = 1 ]; then python bench.py --gpus 1 "$@" > gpurun_out/scale/$name.json 2> gpurun_out/scale/$name.err
  else $T --nproc-per-node $n --master-port $port bench.py --gpus $n "$@" > gpurun_out/scale/$name.json 2> gpurun_out/scale/$name.err; fi
  python - "$name" <<'PY'
import json, sys
name = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/scale/%s.json" % name).read().strip().splitlines()[-1])
    print("%-28s N=%d  %10.0f img/s  %.4f ms/step  e2e %.4f ms  exposed %s  ok=%s" % (name, d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"],
          d.get("exposed_comm_ms_per_step"), d.get("cross_rank_ok")))
except Exception as e:
    print(name, "FAILED", e)
PY
}
for n in 1 2 4 8; do run convnet_ours_w${n}_20 $n --steps 20 --warmup 5; done
run convnet_ours_w8_200 8 --steps 200 --warmup 20
run convnet_ours_w1_200 1 --steps 200 --warmup 20
for n in 1 2 4 8; do run convnet_ref_w${n} $n --impl reference --steps 200 --warmup 20; done
run convnet_sixcall_w8 8 --six-call --steps 200 --warmup 20
run convnet_b4096_ours_w8 8 --batch 4096 --steps 30 --warmup 5
run convnet_b4096_ref_w8 8 --impl reference --batch 4096 --steps 30 --warmup 5
run convnet_b4096_ours_w1 1 --batch 4096 --steps 30 --warmup 5
run r18_ours_w1 1 --model resnet18 --steps 50 --warmup 5
run r18_ours_w2 2 --model resnet18 --steps 50 --warmup 5
run r18_ours_w8 8 --model resnet18 --steps 50 --warmup 5
run r18_ref_w2 2 --impl reference --model resnet18 --steps 50 --warmup 5
run r18_ref_w8 8 --impl reference --model resnet18 --steps 50 --warmup 5
run r50_ours_w1 1 --model resnet50 --steps 20 --warmup 5
run r50_ours_w8 8 --model resnet50 --steps 20 --warmup 5
