# Convenience targets; everything they run is documented in README.md / INTEGRATION.md.
PY ?= python

.PHONY: build test gpu-test bench bench-reference smoke sass clean

build:            ## g++ (host runtime) + nvcc sm_100a (kernels) -> pytorch_distributed_learning_b200/_native/*.so
	$(PY) -m pytorch_distributed_learning_b200._build

test: build       ## CPU tests (multi-process paths run on the C++ TCP backend)
	$(PY) -m pytest tests -x -q -m "not gpu"

gpu-test: build   ## needs a B200; multi-GPU tests skip themselves on a single-GPU box
	$(PY) -m pytest tests -x -q -m gpu

smoke: build      ## one tiny training step of both model families on cuda:0
	$(PY) -c "import __graft_entry__ as g; g.smoke()"

bench: build      ## headline benchmark, 1 GPU (N GPUs: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N)
	$(PY) bench.py

bench-reference:  ## the unmodified reference on stock torch DDP + NCCL, same metric and config
	$(PY) bench.py --impl reference

sass:             ## regenerate profiles/sass_evidence.md from the built kernel library (no GPU needed)
	$(PY) scripts/sass_evidence.py

clean:
	rm -f pytorch_distributed_learning_b200/_native/*.so
	rm -rf build .pytest_cache
