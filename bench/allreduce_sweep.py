#!/usr/bin/env python
"""Allreduce bandwidth sweep (BASELINE config 5): every peer-memory variant vs NCCL, 1 KB .. 256 MB.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench/allreduce_sweep.py [--max-mb 256]

Buffers live in one symmetric segment (in | out | ring staging), kernels are launched back to back
on one stream and timed with CUDA events (max over ranks); busbw = S/t * 2(W-1)/W (NCCL convention),
reported against the 900 GB/s per-direction NVLink 5 nominal and the 770 GB/s measured peer-copy figure.
NCCL (the reference's data plane, launch_dist.py:49) is timed the same way on the same sizes.
Rank 0 prints one JSON line per (size, algorithm) and a markdown table at the end.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import pytorch_distributed_learning_b200 as pdl  # noqa: E402
from pytorch_distributed_learning_b200 import _lib  # noqa: E402
from pytorch_distributed_learning_b200.parallel.b200_backend import ALGOS, _ptr_array  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=256)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--no-nccl", action="store_true")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", os.environ["RANK"]))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = pdl.distributed
    dist.init_process_group(backend="nccl")
    pg = dist._default()
    pg._ensure_cuda(local)
    L = _lib.cuda()
    max_bytes = args.max_mb << 20
    pad = (L.pdl_signal_pad_bytes() + 4095) // 4096 * 4096
    seg = pg.new_segment(pad + 3 * max_bytes)
    state = pg.new_comm_state()
    pads = _ptr_array([seg.ptr(r) for r in range(world)])
    ins = _ptr_array([seg.ptr(r, pad) for r in range(world)])
    outs = _ptr_array([seg.ptr(r, pad + max_bytes) for r in range(world)])
    rings = _ptr_array([seg.ptr(r, pad + 2 * max_bytes) for r in range(world)])
    mc_in = ctypes.c_void_p(seg.mc_ptr + pad) if seg.mc_ptr else None
    mc_out = ctypes.c_void_p(seg.mc_ptr + pad + max_bytes) if seg.mc_ptr else None
    x = seg.tensor(pad, max_bytes // 4)
    y = seg.tensor(pad + max_bytes, max_bytes // 4)
    x.copy_(torch.randn(max_bytes // 4, device=dev))
    stream = torch.cuda.current_stream()

    nccl_ok = False
    if not args.no_nccl and world > 1:
        import torch.distributed as tdist

        # under torch.distributed.run both rendezvous ride on the agent's store at MASTER_PORT (distinct keys)
        tdist.init_process_group(backend="nccl", device_id=dev)
        nccl_ok = True
        nccl_buf = torch.randn(max_bytes // 4, device=dev)

    def timed(fn, iters):
        """Device time per call: `iters` back-to-back launches captured in ONE CUDA graph (no Python/launch gaps)."""
        fn()
        torch.cuda.synchronize()
        pg._ctl.barrier()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(iters):
                fn()
        g.replay()
        torch.cuda.synchronize()
        pg._ctl.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.replay()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / iters
        return max(pg.allgather_object(ms))

    sizes = []
    s = 1024
    while s <= max_bytes:
        sizes.append(s)
        s *= 4
    if sizes[-1] != max_bytes:
        sizes.append(max_bytes)
    names = ["oneshot", "twoshot", "ring"] + (["nvls"] if seg.mc_ptr else [])
    results = []
    for nbytes in sizes:
        n4 = nbytes // 16
        row = {"bytes": nbytes}
        iters = args.iters if nbytes <= (16 << 20) else max(5, args.iters // 4)
        for name in names:
            algo = ALGOS[name]
            if name == "oneshot" and nbytes > (64 << 20):
                continue
            blocks = L.pdl_allreduce_blocks(n4, world, algo)

            def launch(algo=algo, blocks=blocks):
                rc = L.pdl_allreduce_launch(algo, rank, world, 0, n4, ins, outs, rings, pads, mc_in, mc_out, None, None, None,
                                            ctypes.c_void_p(state.data_ptr()), blocks, 1.0 / world, 30.0, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
                assert rc == 0, L.pdl_last_error()

            ms = timed(launch, iters)
            row[name] = ms
        if nccl_ok:
            view = nccl_buf[: nbytes // 4]
            row["nccl"] = timed(lambda: tdist.all_reduce(view), iters)
        results.append(row)
        if rank == 0:
            f = 2.0 * (world - 1) / world
            out = {"bytes": nbytes, "world": world}
            for k, v in row.items():
                if k != "bytes":
                    out[k + "_us"] = round(v * 1e3, 2)
                    out[k + "_busbw_GBs"] = round(nbytes / (v * 1e-3) * f / 1e9, 2)
            print(json.dumps(out), flush=True)
    err = int(state.view(torch.int32)[-1].item())
    if rank == 0:
        f = 2.0 * (world - 1) / world
        cols = names + (["nccl"] if nccl_ok else [])
        print("\n| bytes | " + " | ".join("%s us (busbw GB/s)" % c for c in cols) + " |")
        print("|---|" + "---|" * len(cols))
        for row in results:
            cells = []
            for c in cols:
                if c in row:
                    cells.append("%.1f (%.0f)" % (row[c] * 1e3, row["bytes"] / (row[c] * 1e-3) * f / 1e9))
                else:
                    cells.append("-")
            print("| %d | %s |" % (row["bytes"], " | ".join(cells)))
        print("\nmulticast (NVLS) available: %s; peer-wait timeouts: %d; world %d" % (bool(seg.mc_ptr), err, world))
    if nccl_ok:
        tdist.destroy_process_group()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
