"""Rendezvous key-value store (Python face of csrc/host/store.cpp).

``TCPStore(host, port, world_size, is_master)``: rank 0 starts the C++ server in-process and
every rank (rank 0 included) talks to it through a client socket.  ``PrefixStore`` namespaces
keys per process group, as the reference's stack does with ``PrefixStore("default_pg", ...)``
(reference call sites: mpspawn_dist.py:49-54, example_mp.py:37-42).
"""
from __future__ import annotations

import ctypes
from ctypes import byref, c_char_p, c_int, c_int64, c_uint64, c_void_p

from .. import _lib

DEFAULT_TIMEOUT_S = 1800.0


class StoreTimeout(RuntimeError):
    pass


class TCPStore:
    def __init__(self, host, port, world_size=None, is_master=False, timeout=DEFAULT_TIMEOUT_S):
        L = _lib.host()
        self.host, self.world_size, self.is_master, self.timeout = host, world_size, is_master, float(timeout)
        self._server = None
        self._client = None
        if is_master:
            bound = c_int(0)
            self._server = L.pdl_store_server_start(host.encode(), int(port), byref(bound))
            if not self._server:
                raise RuntimeError("store: cannot listen on %s:%s (address already in use?)" % (host, port))
            port = bound.value
        self.port = int(port)
        connect_host = "127.0.0.1" if is_master else host
        self._client = L.pdl_store_client_connect(connect_host.encode(), self.port, self.timeout)
        if not self._client:
            self.close()
            raise StoreTimeout("store: cannot reach the rendezvous server at %s:%d within %.0f s" % (host, self.port, self.timeout))

    # -- key/value API ---------------------------------------------------------------
    def set(self, key, value):
        if isinstance(value, str):
            value = value.encode()
        value = bytes(value)
        if _lib.host().pdl_store_set(self._client, key.encode(), value, len(value)) != 0:
            raise RuntimeError("store: connection lost during set(%r)" % key)

    def get(self, key, timeout=None):
        out, n = c_void_p(), c_uint64()
        rc = _lib.host().pdl_store_get(self._client, key.encode(), float(self.timeout if timeout is None else timeout), byref(out), byref(n))
        if rc == 1:
            raise StoreTimeout("store: timed out waiting for key %r" % key)
        if rc != 0:
            raise RuntimeError("store: connection lost during get(%r)" % key)
        data = ctypes.string_at(out, n.value)
        _lib.host().pdl_free(out)
        return data

    def add(self, key, amount):
        res = c_int64()
        if _lib.host().pdl_store_add(self._client, key.encode(), int(amount), byref(res)) != 0:
            raise RuntimeError("store: connection lost during add(%r)" % key)
        return res.value

    def _keys(self, keys):
        arr = (c_char_p * len(keys))(*[k.encode() for k in keys])
        return arr

    def wait(self, keys, timeout=None):
        rc = _lib.host().pdl_store_wait(self._client, self._keys(keys), len(keys), float(self.timeout if timeout is None else timeout))
        if rc == 1:
            raise StoreTimeout("store: timed out waiting for keys %r" % (keys,))
        if rc != 0:
            raise RuntimeError("store: connection lost during wait")

    def check(self, keys):
        return _lib.host().pdl_store_check(self._client, self._keys(keys), len(keys)) == 0

    def delete_key(self, key):
        return _lib.host().pdl_store_delete(self._client, key.encode()) == 1

    def num_keys(self):
        return _lib.host().pdl_store_num_keys(self._client)

    def compare_set(self, key, expected, desired):
        expected = expected.encode() if isinstance(expected, str) else bytes(expected)
        desired = desired.encode() if isinstance(desired, str) else bytes(desired)
        out, n = c_void_p(), c_uint64()
        if _lib.host().pdl_store_compare_set(self._client, key.encode(), expected, len(expected), desired, len(desired), byref(out), byref(n)) != 0:
            raise RuntimeError("store: connection lost during compare_set(%r)" % key)
        data = ctypes.string_at(out, n.value)
        _lib.host().pdl_free(out)
        return data

    @property
    def client_handle(self):
        return self._client

    def close(self):
        L = _lib.host()
        if self._client:
            L.pdl_store_client_close(self._client)
            self._client = None
        if self._server:
            L.pdl_store_server_stop(self._server)
            self._server = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PrefixStore:
    """Namespaces every key with ``prefix/`` on top of another store."""

    def __init__(self, prefix, store):
        self.prefix, self.base = prefix, store

    def _k(self, key):
        return self.prefix + "/" + key

    def set(self, key, value):
        self.base.set(self._k(key), value)

    def get(self, key, timeout=None):
        return self.base.get(self._k(key), timeout)

    def add(self, key, amount):
        return self.base.add(self._k(key), amount)

    def wait(self, keys, timeout=None):
        self.base.wait([self._k(k) for k in keys], timeout)

    def check(self, keys):
        return self.base.check([self._k(k) for k in keys])

    def delete_key(self, key):
        return self.base.delete_key(self._k(key))

    def compare_set(self, key, expected, desired):
        return self.base.compare_set(self._k(key), expected, desired)

    @property
    def client_handle(self):
        return self.base.client_handle

    @property
    def full_prefix(self):
        inner = getattr(self.base, "full_prefix", "")
        return (inner + "/" if inner else "") + self.prefix

    @property
    def timeout(self):
        return self.base.timeout
