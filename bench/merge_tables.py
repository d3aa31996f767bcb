#!/usr/bin/env python
"""Merge per-world tables written by ``allreduce_sweep.py --emit-table`` into ``parallel/allreduce_table.json``.

    python bench/merge_tables.py gpurun_out/table_w2.json gpurun_out/table_w4.json gpurun_out/table_w8.json

Adjacent size classes that picked the same (variant, CTA count) pair are folded into one row."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEST = os.path.join(ROOT, "pytorch_distributed_learning_b200", "parallel", "allreduce_table.json")


def fold(rows):
    out = []
    for r in rows:
        key = (r["algo"], r["blocks"], r["algo_no_mc"], r["blocks_no_mc"])
        if out and out[-1]["_key"] == key:
            out[-1]["max_bytes"] = r["max_bytes"]
        else:
            out.append({"max_bytes": r["max_bytes"], "algo": r["algo"], "blocks": r["blocks"], "algo_no_mc": r["algo_no_mc"],
                        "blocks_no_mc": r["blocks_no_mc"], "_key": key})
    for r in out:
        del r["_key"]
    return out


def main():
    worlds = {}
    for path in sys.argv[1:]:
        with open(path) as f:
            for w, rows in json.load(f)["worlds"].items():
                worlds[w] = fold(rows)
    if os.path.exists(DEST):
        with open(DEST) as f:
            old = json.load(f)["worlds"]
        for w, rows in old.items():
            worlds.setdefault(w, rows)
    with open(DEST, "w") as f:
        json.dump({"source": "bench/allreduce_sweep.py --emit-table on 8 x B200 (NVSwitch); fastest (variant, CTAs) per size class, "
                             "see profiles/allreduce_sweep.md", "worlds": dict(sorted(worlds.items(), key=lambda kv: int(kv[0])))}, f, indent=1)
    print("wrote", DEST)


if __name__ == "__main__":
    main()
