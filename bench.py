#!/usr/bin/env python
"""Headline benchmark: images/sec of the reference's DDP training step on N B200s of one node.

    python bench.py --gpus 1 --steps 200 --warmup 20
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P bench.py --gpus 8 ...
    python bench.py --impl reference ...        # the unmodified reference model on stock torch DDP + NCCL

Config (BASELINE.json): MNIST-shaped ConvNet (reference launch_dist.py), synthetic 1x28x28 data,
batch 100 per rank, SGD lr 1e-4, fp32 storage.  ``--model resnet50`` / ``--batch`` select the other
BASELINE configs.  Timing: W warm-up steps, then exactly K steps between barrier+synchronize,
CUDA events on the launching stream, max over ranks.  Inputs rotate through a pool of distinct
batches larger than the 126 MB L2, so no step finds its batch in cache.  ``value`` is the
device-resident-input number; ``e2e`` repeats it through the public API with a pinned-host H2D
copy of each batch and a D2H read of each loss inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

L2_BYTES = 126 * 1024 * 1024


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=200)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--impl", choices=["ours", "reference"], default="ours")
    p.add_argument("--model", choices=["convnet", "resnet18", "resnet50"], default="convnet")
    p.add_argument("--batch", type=int, default=None, help="per-rank batch (default: 100 ConvNet, 128 ResNet-18, 64 ResNet-50)")
    p.add_argument("--eager", action="store_true", help="ours: run the reference-style eager loop instead of the graphed TrainStep")
    p.add_argument("--six-call", action="store_true", help="ours: the reference's loop body verbatim (model(x); criterion; zero_grad; backward; step) "
                   "on this package's classes -- DistributedDataParallel defers the forward and replays the whole-iteration graph in backward()")
    p.add_argument("--e2e-sync", action="store_true", help="end-to-end loop: block on loss.item() after every step (default: read it one step late)")
    p.add_argument("--algorithm", default=None, help="ours: force an allreduce variant (oneshot|twoshot|ring|nvls)")
    p.add_argument("--local-rank", "--local_rank", type=int, default=None)
    return p.parse_args()


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


class ClockSampler:
    """Samples SM clock + throttle reasons of one GPU while the timed region runs (NVML every 5 ms in a thread;
    nvidia-smi polling as the fallback)."""

    REASONS = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, index):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self.thread = None
        self.proc = None

    def start(self):
        try:
            import pynvml

            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))

            def poll():
                while not self._stop.is_set():
                    try:
                        self.samples.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                        mask = int(pynvml.nvmlDeviceGetCurrentClocksEventReasons(h))
                        for name, bit in self.REASONS.items():
                            if mask & bit:
                                self.reasons.add(name)
                    except Exception:
                        pass
                    self._stop.wait(0.001)

            self.thread = threading.Thread(target=poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.thread = None
        try:
            q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                 "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)

            def read():
                names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
                for line in self.proc.stdout:
                    r = [c.strip() for c in line.split(",")]
                    try:
                        self.samples.append(float(r[0]))
                        self.max_mhz = float(r[1])
                    except (ValueError, IndexError):
                        continue
                    for name, v in zip(names, r[2:6]):
                        if v.lower().startswith("active"):
                            self.reasons.add(name)

            self.thread = threading.Thread(target=read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def reset(self):
        """Forget what was sampled so far (warm-up): the record describes the timed region only."""
        self.samples, self.reasons = [], set()

    def stop(self):
        self._stop.set()
        if self.proc is not None:
            time.sleep(0.05)
            self.proc.terminate()
        if self.thread is not None:
            self.thread.join(1.0)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["no clock sample available"], "samples": 0}
        sm = sorted(self.samples)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(sm)}


def ensure_reference_tree():
    """baseline/_ref holds an unmodified copy of the reference (it has no setup.py, so `pip install
    --target baseline/_ref /root/reference` cannot work -- see DESIGN.md); copy the scripts verbatim."""
    dst = os.path.join(ROOT, "baseline", "_ref")
    src = "/root/reference"
    if not os.path.exists(os.path.join(dst, "launch_dist.py")) and os.path.isdir(src):
        import shutil

        os.makedirs(dst, exist_ok=True)
        for f in os.listdir(src):
            if os.path.isfile(os.path.join(src, f)):
                shutil.copy2(os.path.join(src, f), os.path.join(dst, f))
    return dst if os.path.exists(os.path.join(dst, "launch_dist.py")) else None


def cross_rank_check(torch, pg, ddp, step_fn, images, labels, rank, world):
    """What DDP promises (reference README.md:168: every replica holds the same model), checked on the real peers:
    (a) every rank's parameters are bit-identical after the timed training; (b) for one extra step the raw per-rank
    gradients are gathered over the CPU mesh, averaged in fp64 on the host, and compared with what the NVLink allreduce
    kernel left in the averaged arena -- which must also be bit-identical on every rank; (c) no kernel gave up waiting
    for a peer.  Large models are checked on a strided sample of <= 1M elements per rank."""
    import hashlib

    import numpy as np

    torch.cuda.synchronize()

    def digest():
        h = hashlib.sha1()
        for p in ddp.module.parameters():
            h.update(p.detach().cpu().contiguous().numpy().tobytes())
        return h.hexdigest()

    before = pg.allgather_object(digest())
    step_fn(images, labels)
    torch.cuda.synchronize()
    backend = ddp._backend
    total = int(ddp._total)
    stride = max(1, total // (1 << 20))
    raw = backend.g_in[:total:stride].detach().cpu().numpy()
    avg = backend.g_avg[:total:stride].detach().cpu().numpy()
    raws = pg.allgather_object(raw)
    avgs = pg.allgather_object(avg)
    after = pg.allgather_object(digest())
    oracle = np.mean(np.stack([r.astype(np.float64) for r in raws]), axis=0)
    err = max(float(np.max(np.abs(a.astype(np.float64) - oracle))) for a in avgs)
    scale = float(np.max(np.abs(oracle))) or 1.0
    avg_same = all(a.tobytes() == avgs[0].tobytes() for a in avgs)
    moved = after[0] != before[0]  # the extra step really trained
    timeouts = sum(pg.allgather_object(int(pg.check_device_errors(raise_on_error=False))))
    identical = len(set(before)) == 1 and len(set(after)) == 1
    ok = bool(identical and avg_same and moved and timeouts == 0 and err <= 1e-5 * scale + 1e-9)
    return {"replicas_identical": bool(identical), "averaged_grads_identical": bool(avg_same), "allreduce_max_abs_err": err,
            "allreduce_grad_scale": scale, "allreduce_checked_elems": int(raw.size), "peer_timeouts": int(timeouts), "cross_rank_ok": ok}


def main():
    args = parse_args()
    if args.impl == "reference":
        ref_dir = ensure_reference_tree()
        if ref_dir is None:
            print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref is empty and /root/reference is not mounted on this box"}))
            return 0
    import torch

    if not torch.cuda.is_available():
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device visible"}))
        return 0
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    if "MASTER_ADDR" not in os.environ:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(free_port())
        os.environ["RANK"], os.environ["WORLD_SIZE"], os.environ["LOCAL_RANK"] = str(rank), str(world), str(local_rank)
    if world != args.gpus and rank == 0:
        sys.stderr.write("bench: --gpus %d but WORLD_SIZE=%d (launch with torch.distributed.run for N>1); using WORLD_SIZE\n" % (args.gpus, world))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    batch = args.batch or {"convnet": 100, "resnet18": 128, "resnet50": 64}[args.model]
    shape = {"convnet": (1, 28, 28), "resnet18": (3, 32, 32), "resnet50": (3, 224, 224)}[args.model]
    classes = {"convnet": 10, "resnet18": 10, "resnet50": 1000}[args.model]
    opt_kw = dict(lr=1e-4) if args.model == "convnet" else dict(lr=0.02, momentum=0.9, weight_decay=1e-4, nesterov=True)

    # ---------------- model + DDP + step function ----------------
    kernels_per_step = 0
    if args.impl == "reference":
        import torch.distributed as tdist
        import torch.nn as tnn

        sys.path.insert(0, ref_dir)
        tdist.init_process_group(backend="nccl", device_id=dev)
        torch.manual_seed(0)
        if args.model == "convnet":
            import launch_dist as ref_script  # the reference file itself; its __main__ block is not executed

            model = ref_script.ConvNet().to(dev)
        else:
            import torchvision

            model = getattr(torchvision.models, args.model)(num_classes=classes).to(dev)
        criterion = tnn.CrossEntropyLoss().cuda(local_rank)
        optimizer = torch.optim.SGD(model.parameters(), **opt_kw)
        model = tnn.parallel.DistributedDataParallel(model, device_ids=[local_rank], output_device=local_rank)

        def step_fn(images, labels):  # loop body of launch_dist.py:82-92
            images = images.cuda(non_blocking=True)
            labels = labels.cuda(non_blocking=True)
            outputs = model(images)
            loss = criterion(outputs, labels)
            optimizer.zero_grad()
            loss.backward()
            optimizer.step()
            return loss

        def barrier():
            tdist.barrier()

        _sync_word = torch.zeros(1, device=dev)

        def device_barrier():  # NCCL allreduce on the compute stream: a rendezvous on the device, no host round trip
            if world > 1:
                tdist.all_reduce(_sync_word)

        def gather_max(v):
            t = torch.tensor([v], dtype=torch.float64, device=dev)
            tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
            return float(t.item())

        def shutdown():
            tdist.destroy_process_group()
        parallelism, mode = "dp%d (torch DDP + NCCL)" % world, "eager"
    else:
        import pytorch_distributed_learning_b200 as pdl

        dist = pdl.distributed
        dist.init_process_group(backend="nccl")
        torch.manual_seed(0)
        model = (pdl.models.ConvNet() if args.model == "convnet" else getattr(pdl.models, args.model)(num_classes=classes)).to(dev)
        criterion = pdl.nn.CrossEntropyLoss().cuda(local_rank)
        optimizer = pdl.optim.SGD(model.parameters(), **opt_kw)
        model = pdl.nn.parallel.DistributedDataParallel(model, device_ids=[local_rank], output_device=local_rank, algorithm=args.algorithm)
        train_step = pdl.TrainStep(model, criterion, optimizer, use_graph=not args.eager)
        step_fn = train_step
        if args.six_call:
            def step_fn(images, labels):  # loop body of launch_dist.py:82-92, unchanged
                images = images.cuda(non_blocking=True)
                labels = labels.cuda(non_blocking=True)
                outputs = model(images)
                loss = criterion(outputs, labels)
                optimizer.zero_grad()
                loss.backward()
                optimizer.step()
                return loss

        def barrier():
            torch.cuda.synchronize()
            dist.barrier()

        def device_barrier():  # flag kernel over peer memory on the compute stream (B200ProcessGroup.device_barrier)
            if world > 1:
                dist._default().device_barrier()

        def gather_max(v):
            return max(dist._default().allgather_object(float(v)))

        def shutdown():
            dist.destroy_process_group()
        parallelism = "dp%d (P2P NVLink allreduce fused with SGD)" % world
        mode = "cuda-graph" if train_step.graphed else "eager"
        if args.six_call:
            mode = "six-call loop (deferred forward, graph replayed by loss.backward())"

    # ---------------- synthetic data: pool larger than L2 ----------------
    per_batch = batch * (shape[0] * shape[1] * shape[2]) * 4
    pool = max(4, min(1024, L2_BYTES // per_batch + 2))
    g = torch.Generator().manual_seed(1234 + rank)
    host_x = [torch.randn((batch,) + shape, generator=g).pin_memory() for _ in range(pool)]
    host_y = [torch.randint(0, classes, (batch,), generator=g).pin_memory() for _ in range(pool)]
    dev_pool = min(pool, max(4, (8 << 30) // per_batch))
    dev_x = [t.to(dev) for t in host_x[:dev_pool]]
    dev_y = [t.to(dev) for t in host_y[:dev_pool]]

    # ---------------- device-timed run (inputs resident on the GPU) ----------------
    # Order matters for the 20 x 0.12 ms regions the driver times: everything slow or rank-skewed on the host (NVML
    # start-up, thread creation, event allocation) happens BEFORE the rendezvous; the last thing in front of
    # `start.record()` is a barrier executed ON THE DEVICE (ours: a flag kernel over peer memory; reference: NCCL's
    # barrier allreduce), so every GPU enters the timed region within microseconds of its peers and a late host only
    # delays its own queue, never another rank's clock.
    sampler = ClockSampler(local_rank)
    sampler.start()
    for i in range(max(3, args.warmup)):
        step_fn(dev_x[i % dev_pool], dev_y[i % dev_pool])
    if args.impl == "ours":
        kernels_per_step = train_step.kernels_per_step() or 0
        if args.six_call:
            engines = [e for e, _ in getattr(model, "_graph_engines", {}).values() if e is not None]
            kernels_per_step = engines[0].kernels_per_step() if engines else 0
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    barrier()
    torch.cuda.synchronize()
    sampler.reset()
    device_barrier()
    t_wall0 = time.perf_counter()
    start.record()
    for i in range(args.steps):
        step_fn(dev_x[i % dev_pool], dev_y[i % dev_pool])
        marks[i].record()
    stop.record()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    clocks = sampler.stop()
    barrier()
    dev_ms = start.elapsed_time(stop)
    step_ms = [(start if i == 0 else marks[i - 1]).elapsed_time(marks[i]) for i in range(args.steps)]
    dev_ms = gather_max(dev_ms)
    wall_ms = gather_max(wall_ms)
    step_ms = [gather_max(v) for v in step_ms] if args.steps <= 64 else step_ms
    ordered = sorted(step_ms)
    step_stats = {"p50": ordered[len(ordered) // 2], "p90": ordered[min(len(ordered) - 1, int(len(ordered) * 0.9))], "max": ordered[-1],
                  "first5": [round(v, 4) for v in step_ms[:5]]}

    # ---------------- end to end: pinned host batch -> H2D -> step -> loss D2H, every step ----------------
    # Both arms read every step's loss on the host ONE STEP LATE (step i's value is fetched after step i+1 has been
    # issued), so neither arm drains the GPU per step; the last value is fetched before the closing event.  Reference:
    # loss.item() on the previous step's tensor.  Ours: TrainStep.step_async -> PendingLoss (pinned-host ring) with the
    # batch's H2D copy on a copy stream.  --e2e-sync restores a blocking .item() after every step in both arms.
    if args.impl == "ours" and not args.e2e_sync and not args.six_call:
        e2e_step = train_step.step_async
    else:
        e2e_step = step_fn
    for i in range(3):
        float(e2e_step(host_x[i % pool], host_y[i % pool]).item())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    device_barrier()
    e0.record()
    loss_sum = 0.0
    prev = None
    for i in range(args.steps):
        cur = e2e_step(host_x[i % pool], host_y[i % pool])
        if args.e2e_sync:
            loss_sum += float(cur.item())
            continue
        if prev is not None:
            loss_sum += float(prev.item())
        prev = cur
    if prev is not None:
        loss_sum += float(prev.item())
    e1.record()
    torch.cuda.synchronize()
    barrier()
    e2e_ms = gather_max(e0.elapsed_time(e1))

    # ---------------- exposed communication: the same step with the gradient exchange removed ----------------
    # (reference: DDP.no_sync(); ours: an identical model stepped without the DDP wrapper).  Device-timed like the
    # headline; exposed comm = headline ms/step - compute-only ms/step.
    local_ms = None
    if world > 1:
        if args.impl == "reference":
            def local_step(images, labels):
                with model.no_sync():
                    return step_fn(images, labels)
        else:
            torch.manual_seed(0)
            twin = (pdl.models.ConvNet() if args.model == "convnet" else getattr(pdl.models, args.model)(num_classes=classes)).to(dev)
            local_step = pdl.TrainStep(twin, criterion, pdl.optim.SGD(twin.parameters(), **opt_kw), use_graph=not args.eager)
        for i in range(max(3, args.warmup)):
            local_step(dev_x[i % dev_pool], dev_y[i % dev_pool])
        l0, l1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        torch.cuda.synchronize()
        device_barrier()
        l0.record()
        for i in range(args.steps):
            local_step(dev_x[i % dev_pool], dev_y[i % dev_pool])
        l1.record()
        torch.cuda.synchronize()
        barrier()
        local_ms = gather_max(l0.elapsed_time(l1))

    # ---------------- cross-rank correctness (ours, N > 1; outside every timed region) ----------------
    check = None
    if args.impl == "ours" and world > 1:
        check = cross_rank_check(torch, dist._default(), model, step_fn, dev_x[0], dev_y[0], rank, world)

    stats = {}
    if args.impl == "ours":
        try:
            stats = model.get_step_stats()
        except Exception:
            stats = {}
    if rank == 0:
        global_batch = batch * world
        out = {
            "metric": "images/sec", "value": global_batch * args.steps / (dev_ms / 1e3), "unit": "images/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "impl": args.impl,
            "config": {"model": {"convnet": "ConvNet (reference launch_dist.py, MNIST-shaped 1x28x28)", "resnet18": "ResNet-18 (3x32x32, 10 classes)",
                                 "resnet50": "ResNet-50 (3x224x224)"}[args.model],
                       "global_batch": global_batch, "per_gpu_batch": batch, "seq_len": None, "parallelism": parallelism, "optimizer": "SGD %s" % opt_kw,
                       "execution": mode, "l2_policy": "inputs rotate through a pool of %d distinct batches (%.0f MB > 126 MB L2)" % (dev_pool, dev_pool * per_batch / 1e6),
                       "weights": "random init", "timing": "CUDA events, max over ranks; wall %.3f ms/step" % (wall_ms / args.steps)},
            "clocks": clocks,
            "e2e": {"value": global_batch * args.steps / (e2e_ms / 1e3), "unit": "images/s", "ms_per_step": e2e_ms / args.steps,
                    "h2d_bytes_per_step": per_batch + batch * 8, "d2h_bytes_per_step": 4, "mean_loss": loss_sum / args.steps,
                    "loss_read": "blocking .item() after every step" if args.e2e_sync else
                    "every step's loss is read on the host one step late (both arms); all reads inside the timed region"},
            "gpu_launches": int(kernels_per_step * args.steps) if args.impl == "ours" else None,
        }
        if local_ms is not None:
            out["compute_only_ms_per_step"] = local_ms / args.steps
            out["exposed_comm_ms_per_step"] = max(0.0, (dev_ms - local_ms) / args.steps)
        out["step_ms"] = step_stats
        if stats:
            out["config"]["allreduce"] = stats.get("algorithms")
        if check is not None:
            out.update(check)
        print(json.dumps(out))
        sys.stdout.flush()
    shutdown()
    if check is not None and not check["cross_rank_ok"]:
        return 3
    return 0


if __name__ == "__main__":
    sys.exit(main())
