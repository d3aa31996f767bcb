import threading
import time

import pytest

from pytorch_distributed_learning_b200.parallel.store import PrefixStore, StoreTimeout, TCPStore


def test_set_get_add_wait(port):
    master = TCPStore("127.0.0.1", port, 2, is_master=True, timeout=10)
    client = TCPStore("127.0.0.1", port, 2, is_master=False, timeout=10)
    master.set("a", b"\x00\x01binary")
    assert client.get("a") == b"\x00\x01binary"
    assert client.add("ctr", 3) == 3
    assert master.add("ctr", 4) == 7
    assert master.get("ctr") == b"7"
    assert client.check(["a", "ctr"]) and not client.check(["a", "nope"])
    assert client.num_keys() == 2
    assert client.delete_key("a") and not client.delete_key("a")
    with pytest.raises(StoreTimeout):
        client.get("never", timeout=0.2)
    with pytest.raises(StoreTimeout):
        client.wait(["never"], timeout=0.2)
    client.close()
    master.close()


def test_blocking_get_wakes_up(port):
    master = TCPStore("127.0.0.1", port, 2, is_master=True, timeout=10)
    client = TCPStore("127.0.0.1", port, 2, is_master=False, timeout=10)
    got = {}

    def reader():
        got["v"] = client.get("late")

    t = threading.Thread(target=reader)
    t.start()
    time.sleep(0.2)
    master.set("late", "value")
    t.join(5)
    assert got["v"] == b"value"
    client.close()
    master.close()


def test_prefix_and_compare_set(port):
    master = TCPStore("127.0.0.1", port, 1, is_master=True, timeout=10)
    a, b = PrefixStore("A", master), PrefixStore("B", master)
    a.set("k", "1")
    b.set("k", "2")
    assert a.get("k") == b"1" and b.get("k") == b"2"
    assert PrefixStore("inner", a).full_prefix == "A/inner"
    assert master.compare_set("cas", "", "first") == b"first"
    assert master.compare_set("cas", "wrong", "second") == b"first"
    assert master.compare_set("cas", "first", "second") == b"second"
    master.close()


def test_port_in_use(port):
    master = TCPStore("127.0.0.1", port, 1, is_master=True, timeout=5)
    with pytest.raises(RuntimeError):
        TCPStore("127.0.0.1", port, 1, is_master=True, timeout=5)
    master.close()
