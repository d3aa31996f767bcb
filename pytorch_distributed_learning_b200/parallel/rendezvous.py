"""URL rendezvous: ``env://`` and ``tcp://host:port``.

Same contract the reference relies on: ``env://`` reads RANK / WORLD_SIZE / MASTER_ADDR /
MASTER_PORT (explicit ``rank=`` / ``world_size=`` win over the environment; reference
mpspawn_dist.py:49-54 and launch_dist.py:49), ``tcp://`` needs both given explicitly
(reference example_mp.py:37-42).  Rank 0 hosts the store.
"""
from __future__ import annotations

import os
from urllib.parse import parse_qs, urlparse

from .store import DEFAULT_TIMEOUT_S, TCPStore


class RendezvousError(ValueError):
    pass


def _env_error(var):
    return RendezvousError(
        "Error initializing the process group using env:// rendezvous: environment variable %s expected, but not set" % var
    )


def _query(url):
    return {k: v[0] for k, v in parse_qs(urlparse(url).query).items()}


def parse(init_method, rank=-1, world_size=-1):
    """-> (host, port, rank, world_size).  Pure function: easy to unit-test."""
    if init_method is None:
        init_method = "env://"
    u = urlparse(init_method)
    q = _query(init_method)
    if u.scheme == "env":
        if rank is None or rank < 0:
            if "rank" in q:
                rank = int(q["rank"])
            elif "RANK" in os.environ:
                rank = int(os.environ["RANK"])
            else:
                raise _env_error("RANK")
        if world_size is None or world_size < 0:
            if "world_size" in q:
                world_size = int(q["world_size"])
            elif "WORLD_SIZE" in os.environ:
                world_size = int(os.environ["WORLD_SIZE"])
            else:
                raise _env_error("WORLD_SIZE")
        for var in ("MASTER_ADDR", "MASTER_PORT"):
            if var not in os.environ:
                raise _env_error(var)
        host, port = os.environ["MASTER_ADDR"], int(os.environ["MASTER_PORT"])
    elif u.scheme == "tcp":
        if not u.port:
            raise RendezvousError("tcp:// rendezvous: port number missing in %r" % init_method)
        if rank is None or rank < 0:
            if "rank" not in q:
                raise RendezvousError("tcp:// rendezvous: rank parameter missing")
            rank = int(q["rank"])
        if world_size is None or world_size < 0:
            if "world_size" not in q:
                raise RendezvousError("tcp:// rendezvous: world size parameter missing")
            world_size = int(q["world_size"])
        host, port = u.hostname, int(u.port)
    else:
        raise RendezvousError("No rendezvous handler for %s:// (supported: env://, tcp://)" % u.scheme)
    if not (0 <= rank < world_size):
        raise RendezvousError("rank %d is outside [0, world_size=%d)" % (rank, world_size))
    return host, port, rank, world_size


_agent_rendezvous_count = 0


def _port_via_agent_store(host, port, rank, world_size, timeout):
    """Under ``python -m torch.distributed.run`` the elastic agent itself listens on MASTER_PORT
    (TORCHELASTIC_USE_AGENT_STORE=True), so rank 0 cannot host this package's store there.  The agent's
    store is used for exactly one thing: rank 0 starts our C++ store on an ephemeral port and publishes
    the number; everything else (handles, barriers, endpoints) goes through our own store."""
    global _agent_rendezvous_count
    from datetime import timedelta

    from torch.distributed import TCPStore as AgentStore

    agent = AgentStore(host, port, world_size, is_master=False, timeout=timedelta(seconds=min(timeout, 600.0)))
    key = "pdl_b200/store_port/%s/%d" % (os.environ.get("TORCHELASTIC_RESTART_COUNT", "0"), _agent_rendezvous_count)
    _agent_rendezvous_count += 1
    if rank == 0:
        store = TCPStore(host, 0, world_size, is_master=True, timeout=timeout)
        agent.set(key, str(store.port))
        return store
    real_port = int(agent.get(key).decode())
    return TCPStore(host, real_port, world_size, is_master=False, timeout=timeout)


def rendezvous(init_method, rank=-1, world_size=-1, timeout=DEFAULT_TIMEOUT_S):
    host, port, rank, world_size = parse(init_method, rank, world_size)
    scheme = urlparse(init_method or "env://").scheme
    if scheme == "env" and os.environ.get("TORCHELASTIC_USE_AGENT_STORE", "") == "True":
        store = _port_via_agent_store(host, port, rank, world_size, timeout)
    else:
        store = TCPStore(host, port, world_size, is_master=(rank == 0), timeout=timeout)
    return store, rank, world_size
