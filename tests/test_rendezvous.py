import pytest

from pytorch_distributed_learning_b200.parallel.rendezvous import RendezvousError, parse


def test_env(monkeypatch):
    monkeypatch.setenv("MASTER_ADDR", "10.0.0.1")
    monkeypatch.setenv("MASTER_PORT", "1234")
    monkeypatch.setenv("RANK", "3")
    monkeypatch.setenv("WORLD_SIZE", "8")
    assert parse("env://") == ("10.0.0.1", 1234, 3, 8)
    assert parse(None) == ("10.0.0.1", 1234, 3, 8)
    # explicit arguments win over the environment (mpspawn_dist.py:49-54 passes both)
    assert parse("env://", rank=1, world_size=2) == ("10.0.0.1", 1234, 1, 2)
    assert parse("env://?rank=5&world_size=6") == ("10.0.0.1", 1234, 5, 6)


@pytest.mark.parametrize("missing", ["RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"])
def test_env_missing(monkeypatch, missing):
    for k, v in {"MASTER_ADDR": "h", "MASTER_PORT": "1", "RANK": "0", "WORLD_SIZE": "1"}.items():
        monkeypatch.setenv(k, v)
    monkeypatch.delenv(missing)
    with pytest.raises(RendezvousError, match=missing):
        parse("env://")


def test_tcp():
    assert parse("tcp://10.157.106.151:12345", rank=1, world_size=4) == ("10.157.106.151", 12345, 1, 4)
    assert parse("tcp://h:1?rank=0&world_size=2") == ("h", 1, 0, 2)
    with pytest.raises(RendezvousError, match="rank"):
        parse("tcp://h:1", world_size=2)
    with pytest.raises(RendezvousError, match="world size"):
        parse("tcp://h:1", rank=0)
    with pytest.raises(RendezvousError, match="port"):
        parse("tcp://h", rank=0, world_size=1)


def test_bad_scheme_and_range():
    with pytest.raises(RendezvousError):
        parse("file:///tmp/x", rank=0, world_size=1)
    with pytest.raises(RendezvousError):
        parse("tcp://h:1", rank=2, world_size=2)
