#!/usr/bin/env python
"""Allreduce bandwidth sweep (BASELINE config 5): every peer-memory variant vs NCCL, 1 KB .. 1 GiB.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench/allreduce_sweep.py \
        [--max-mb 1024] [--emit-table out.json] [--markdown out.md]

Buffers live in one symmetric segment (in | out | ring staging) plus the push variant's inbox; `iters` launches are
captured in ONE CUDA graph and timed with CUDA events (max over ranks); busbw = S/t * 2(W-1)/W (NCCL convention),
reported against the 900 GB/s per-direction NVLink 5 nominal and the 770 GB/s measured peer-copy figure.
NCCL (the reference's data plane, launch_dist.py:49) is timed the same way on the same sizes.

Every variant is timed at several CTA counts; the fastest (variant, CTA count) per size -- with and without NVSwitch
multicast -- is what ``--emit-table`` writes in the format of ``parallel/allreduce_table.json`` (merge the per-world
files with ``bench/merge_tables.py``).  Rank 0 prints one JSON line per size and a markdown table at the end.
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

PUSH_MAX = 4 << 20      # the push variant moves (W-1) x S out of every GPU: a small-message algorithm
ONESHOT_MAX = 16 << 20  # same for one-shot's inbound traffic


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--no-nccl", action="store_true")
    ap.add_argument("--emit-table", default=None)
    ap.add_argument("--markdown", default=None)
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
    push_row4 = min(PUSH_MAX, max_bytes) // 16
    inbox = pg.new_segment(int(L.pdl_allreduce_push_inbox_bytes(push_row4, world)), want_multicast=False)
    state = pg.new_comm_state()
    pads = _ptr_array([seg.ptr(r) for r in range(world)])
    ins = _ptr_array([seg.ptr(r, pad) for r in range(world)])
    outs = _ptr_array([seg.ptr(r, pad + max_bytes) for r in range(world)])
    rings = _ptr_array([seg.ptr(r, pad + 2 * max_bytes) for r in range(world)])
    inboxes = _ptr_array([inbox.ptr(r) for r in range(world)])
    mc_in = ctypes.c_void_p(seg.mc_ptr + pad) if seg.mc_ptr else None
    mc_out = ctypes.c_void_p(seg.mc_ptr + pad + max_bytes) if seg.mc_ptr else None
    x = seg.tensor(pad, max_bytes // 4)
    chunk = 64 << 20
    for off in range(0, max_bytes // 4, chunk):  # fill in pieces: no 1 GiB temporary
        n = min(chunk, max_bytes // 4 - off)
        x[off:off + n].copy_(torch.randn(n, device=dev))

    nccl_ok = False
    if not args.no_nccl and world > 1:
        import torch.distributed as tdist

        # under torch.distributed.run both rendezvous ride on the agent's store at MASTER_PORT (distinct keys)
        tdist.init_process_group(backend="nccl", device_id=dev)
        nccl_ok = True
        nccl_buf = torch.empty(max_bytes // 4, device=dev)
        nccl_buf.copy_(x)

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
    names = ["push", "oneshot", "twoshot", "ring"] + (["nvls"] if seg.mc_ptr else [])
    max_blocks = L.pdl_comm_max_blocks()
    results = []
    for nbytes in sizes:
        n4 = nbytes // 16
        row = {"bytes": nbytes, "blocks": {}}
        iters = args.iters if nbytes <= (16 << 20) else max(4, args.iters // 4)
        for name in names:
            algo = ALGOS[name]
            if (name == "oneshot" and nbytes > ONESHOT_MAX) or (name == "push" and nbytes > PUSH_MAX):
                continue
            base = L.pdl_allreduce_blocks(n4, world, algo)
            cands = sorted({base, max(1, base // 2), min(max_blocks, n4, base * 2)}) if nbytes >= (64 << 10) else [base]
            best = None
            for blocks in cands:
                aux = inboxes if name == "push" else rings
                row4 = push_row4 if name == "push" else n4

                def launch(algo=algo, blocks=blocks, aux=aux, row4=row4):
                    rc = L.pdl_allreduce_launch_ex(algo, rank, world, 0, n4, row4, ins, outs, aux, pads, mc_in, mc_out, None, None, None,
                                                   ctypes.c_void_p(state.data_ptr()), blocks, 1.0 / world, 30.0,
                                                   ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
                    assert rc == 0, L.pdl_last_error()

                ms = timed(launch, iters)
                if best is None or ms < best[0]:
                    best = (ms, blocks)
            row[name], row["blocks"][name] = best
        if nccl_ok:
            view = nccl_buf[: nbytes // 4]
            row["nccl"] = timed(lambda: tdist.all_reduce(view), iters)
        results.append(row)
        if rank == 0:
            f = 2.0 * (world - 1) / world
            out = {"bytes": nbytes, "world": world}
            for k, v in row.items():
                if k not in ("bytes", "blocks"):
                    out[k + "_us"] = round(v * 1e3, 2)
                    out[k + "_busbw_GBs"] = round(nbytes / (v * 1e-3) * f / 1e9, 2)
            out["blocks"] = row["blocks"]
            print(json.dumps(out), flush=True)
    err = int(state.view(torch.int32)[L.pdl_comm_state_error_offset() // 4].item())
    if rank == 0:
        f = 2.0 * (world - 1) / world
        cols = names + (["nccl"] if nccl_ok else [])
        lines = ["| bytes | " + " | ".join("%s us (busbw GB/s)" % c for c in cols) + " | best | vs nccl |", "|---|" + "---|" * (len(cols) + 2)]
        table_rows = []
        for row in results:
            cells = []
            for c in cols:
                if c in row:
                    tag = "/%d" % row["blocks"][c] if c in row["blocks"] else ""
                    cells.append("%.1f (%.0f)%s" % (row[c] * 1e3, row["bytes"] / (row[c] * 1e-3) * f / 1e9, tag))
                else:
                    cells.append("n/a")
            ours = {c: row[c] for c in names if c in row}
            best = min(ours, key=ours.get)
            no_mc = {c: v for c, v in ours.items() if c != "nvls"}
            best_no_mc = min(no_mc, key=no_mc.get)
            speed = "%.2fx" % (row["nccl"] / ours[best]) if "nccl" in row else "-"
            lines.append("| %d | %s | %s | %s |" % (row["bytes"], " | ".join(cells), best, speed))
            table_rows.append({"max_bytes": row["bytes"], "algo": best, "blocks": row["blocks"][best], "algo_no_mc": best_no_mc,
                               "blocks_no_mc": row["blocks"][best_no_mc], "us": round(ours[best] * 1e3, 2),
                               "nccl_us": round(row["nccl"] * 1e3, 2) if "nccl" in row else None})
        # size classes: a measured size stands for everything up to the geometric midpoint to the next one
        for i, tr in enumerate(table_rows):
            tr["max_bytes"] = int(tr["max_bytes"] * 2) if i + 1 < len(table_rows) else None
        lines.append("")
        lines.append("cells: microseconds (bus bandwidth GB/s)/CTAs of the fastest CTA count tried; multicast (NVLS) available: %s; "
                     "peer-wait timeouts: %d; world %d" % (bool(seg.mc_ptr), err, world))
        text = "\n".join(lines)
        print("\n" + text)
        if args.markdown:
            with open(args.markdown, "w") as fh:
                fh.write(text + "\n")
        if args.emit_table:
            with open(args.emit_table, "w") as fh:
                json.dump({"worlds": {str(world): table_rows}}, fh, indent=1)
    if nccl_ok:
        tdist.destroy_process_group()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
