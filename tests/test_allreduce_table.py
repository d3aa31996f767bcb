"""The measured allreduce selection table (parallel/allreduce_table.json) and the code that reads it (SURVEY 5.9 "Selection":
per bucket size, from measured bus bandwidth).  CPU-only: the table is data, the lookup is plain Python."""
import json
import os

import pytest

b200 = pytest.importorskip("pytorch_distributed_learning_b200.parallel.b200_backend")

TABLE = os.path.join(os.path.dirname(b200.__file__), "allreduce_table.json")


def test_table_is_well_formed_and_covers_every_size():
    raw = json.load(open(TABLE))
    assert set(raw["worlds"]) >= {"2", "4", "8"}
    for w, rows in raw["worlds"].items():
        bounds = [r["max_bytes"] for r in rows]
        assert bounds[-1] is None, "the last size class of world %s must be open-ended" % w
        finite = bounds[:-1]
        assert finite == sorted(finite) and len(set(finite)) == len(finite)
        for r in rows:
            assert r["algo"] in b200.ALGOS and r["algo_no_mc"] in b200.ALGOS and r["algo_no_mc"] != "nvls"
            for key in ("blocks", "blocks_no_mc"):
                assert r[key] is None or 1 <= r[key] <= 128


def test_lookup_follows_the_table(monkeypatch):
    monkeypatch.delenv("PDL_ALLREDUCE_TABLE", raising=False)
    b200._TABLE = None
    raw = json.load(open(TABLE))["worlds"]
    for w in (2, 4, 8):
        for nbytes in (1, 1024, 454696, 25 << 20, 1 << 30):
            row = b200.table_row(nbytes, w)
            expect = next(r for r in raw[str(w)] if r["max_bytes"] is None or nbytes <= r["max_bytes"])
            assert row == expect
            assert b200.choose_algorithm(nbytes, w, True) == b200.ALGOS[row["algo"]]
            assert b200.choose_algorithm(nbytes, w, False) == b200.ALGOS[row["algo_no_mc"]]
    # a world size that was not measured borrows the next larger one; world 1 never communicates
    assert b200.table_row(4096, 3) == b200.table_row(4096, 4)
    assert b200.table_row(4096, 6) == b200.table_row(4096, 8)
    assert b200.choose_algorithm(4096, 1, True) == b200.ALGOS["oneshot"]


def test_override_and_switches(monkeypatch):
    b200._TABLE = None
    assert b200.choose_algorithm(1 << 20, 8, True, override="ring") == b200.ALGOS["ring"]
    with pytest.raises(RuntimeError, match="multicast"):
        b200.choose_algorithm(1 << 20, 8, False, override="nvls")
    with pytest.raises(ValueError, match="unknown"):
        b200.choose_algorithm(1 << 20, 8, True, override="butterfly")
    monkeypatch.setenv("PDL_ALLREDUCE_TABLE", "off")  # the rule of thumb: NVLS with multicast, one-shot / two-shot without
    assert b200.choose_algorithm(1 << 20, 8, True) == b200.ALGOS["nvls"]
    assert b200.choose_algorithm(1 << 10, 8, False) == b200.ALGOS["oneshot"]
    assert b200.choose_algorithm(1 << 20, 8, False) == b200.ALGOS["twoshot"]


def test_small_messages_take_the_one_round_variant():
    """What the sweep found on every world size: below 256 KiB nothing beats one flag round (push)."""
    b200._TABLE = None
    for w in (2, 4, 8):
        for nbytes in (1024, 3328, 65536):
            assert b200.table_row(nbytes, w)["algo"] in ("push", "twoshot")  # (W=8, 16 KiB: two-shot and push tie within noise)
        assert b200.table_row(3328, w)["algo"] == "push"
