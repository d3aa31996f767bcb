"""utils.tracing: host ranges around DDP forward / bucket launches / TrainStep, exported as Chrome-trace JSON per rank."""
import json

import _workers
from conftest import free_port

import pytorch_distributed_learning_b200 as pdl
from pytorch_distributed_learning_b200.utils import tracing


def test_ranges_are_noops_when_switched_off(monkeypatch):
    monkeypatch.delenv("PDL_TRACE", raising=False)
    monkeypatch.delenv("PDL_NVTX", raising=False)
    tracing.reset()
    before = len(tracing.events())
    with tracing.trace_range("x", a=1):
        pass
    assert not tracing.enabled() and len(tracing.events()) == before
    tracing.reset()


def test_trace_files_per_rank(tmp_path):
    pdl.multiprocessing.spawn(_workers.tracing_worker, args=(2, free_port(), str(tmp_path)), nprocs=2)
    for r in range(2):
        evs = json.load(open(tmp_path / ("trace.rank%d.json" % r)))["traceEvents"]
        names = [e["name"] for e in evs]
        assert names.count("pdl.TrainStep") == 2 and names.count("pdl.DistributedDataParallel.forward") == 2
        buckets = [e for e in evs if e["name"] == "pdl.bucket"]
        assert len(buckets) == 2 and buckets[0]["args"]["params"] == 8  # the ConvNet is one bucket of 8 tensors
        step = [e for e in evs if e["name"] == "pdl.TrainStep"][0]
        inner = [e for e in evs if e["name"] != "pdl.TrainStep" and step["ts"] <= e["ts"] <= step["ts"] + step["dur"]]
        assert inner and all(e["dur"] >= 0 and e["ph"] == "X" for e in evs)
