"""Python face of the C++ TCP store (csrc/store/tcp_store.cpp) — the rendezvous key/value service
(the reference reaches torch's TCPStore through ``dist.init_process_group``,
ref:pytorch/resnet/main.py:148, ref:pytorch/unet/train.py:255, ref:pytorch/hello_world/hello_world.py:34)."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

from .. import _native as N


class TCPStore:
    """Rank 0 hosts (``is_master=True``); everyone, including rank 0, talks to it through a client socket."""

    def __init__(self, host: str, port: int, is_master: bool, timeout_s: float = 300.0):
        self._lib = N.host()
        self._server = None
        self.timeout_ms = int(timeout_s * 1000)
        if is_master:
            out = C.c_int(0)
            self._server = self._lib.dmpi_store_server_start(int(port), C.byref(out))
            if not self._server:
                raise RuntimeError(f"could not bind the rendezvous store on port {port}")
            port = out.value
        self.host, self.port = host, int(port)
        self._client = self._lib.dmpi_store_connect(host.encode(), self.port, self.timeout_ms)
        if not self._client:
            raise TimeoutError(f"could not reach the rendezvous store at {host}:{self.port}")

    def set(self, key: str, value: bytes) -> None:
        buf = C.create_string_buffer(value, len(value))
        if self._lib.dmpi_store_set(self._client, key.encode(), buf, len(value)) != 0:
            raise RuntimeError(f"store.set({key}) failed")

    def get(self, key: str, timeout_s: Optional[float] = None, max_bytes: int = 1 << 20) -> bytes:
        buf = C.create_string_buffer(max_bytes)
        t = self.timeout_ms if timeout_s is None else int(timeout_s * 1000)
        n = self._lib.dmpi_store_get(self._client, key.encode(), buf, max_bytes, t)
        if n == -2:
            return self.get(key, timeout_s, max_bytes * 16)
        if n < 0:
            raise TimeoutError(f"store.get({key}) timed out after {t} ms")
        return buf.raw[:n]

    def add(self, key: str, delta: int) -> int:
        v = self._lib.dmpi_store_add(self._client, key.encode(), int(delta))
        if v == -(1 << 63):
            raise RuntimeError(f"store.add({key}) failed")
        return v

    def check(self, key: str) -> bool:
        return self._lib.dmpi_store_check(self._client, key.encode()) == 1

    # ---- collectives built on the store (control plane only: handles, hashes, barriers)
    def barrier(self, name: str, rank: int, world: int, timeout_s: Optional[float] = None) -> None:
        n = self.add(f"__barrier__/{name}/count", 1)
        if n == world:
            self.set(f"__barrier__/{name}/done", b"1")
        self.get(f"__barrier__/{name}/done", timeout_s)

    def all_gather(self, name: str, rank: int, world: int, payload: bytes) -> List[bytes]:
        self.set(f"__gather__/{name}/{rank}", payload)
        return [self.get(f"__gather__/{name}/{r}") for r in range(world)]

    def close(self) -> None:
        if self._client:
            self._lib.dmpi_store_close(self._client)
            self._client = None
        if self._server:
            self._lib.dmpi_store_server_stop(self._server)
            self._server = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass
