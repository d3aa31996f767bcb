"""``python -m pytorch_distributed_learning_b200.launch`` (see parallel/launch.py)."""
from .parallel.launch import build_parser, main, run, worker_cmd, worker_env  # noqa: F401

if __name__ == "__main__":
    main()
