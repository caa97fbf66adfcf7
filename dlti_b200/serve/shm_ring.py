"""Shared-memory control ring between the TP driver rank and its workers on one node: ctypes
binding of csrc/shm_ring.cpp. One message per engine step (header + packed step metadata)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from ..ops import _lib


class ShmRing:
    def __init__(self, name: str, reader: int, n_readers: int, n_slots: int = 8, slot_bytes: int = 4 << 20):
        """reader < 0: create the ring (driver); reader = 0..n_readers-1: attach (worker `reader`)."""
        lib = _lib.cpu_lib()
        lib.dlti_shm_ring_open.restype = C.c_void_p
        lib.dlti_shm_ring_open.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int]
        lib.dlti_shm_ring_write.restype = C.c_int
        lib.dlti_shm_ring_write.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int]
        lib.dlti_shm_ring_read.restype = C.c_longlong
        lib.dlti_shm_ring_read.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int]
        lib.dlti_shm_ring_close.restype = None
        lib.dlti_shm_ring_close.argtypes = [C.c_void_p]
        self.lib, self.slot_bytes = lib, slot_bytes
        self.h = lib.dlti_shm_ring_open(name.encode(), reader, n_readers, n_slots, slot_bytes)
        if not self.h:
            raise OSError(f"shared-memory ring {name}: open failed (reader={reader})")
        self._buf = np.empty(slot_bytes, dtype=np.uint8)

    def write(self, arr: np.ndarray, timeout_ms: int = 60000) -> None:
        a = np.ascontiguousarray(arr)
        rc = self.lib.dlti_shm_ring_write(self.h, a.ctypes.data, a.nbytes, timeout_ms)
        if rc != 0:
            raise RuntimeError(f"shared-memory ring write failed ({rc}: {'timeout, a worker stopped reading' if rc == -2 else 'message too large'})")

    def read(self, timeout_ms: int = 0) -> np.ndarray:
        """Next message as a uint8 array (a view of an internal buffer: consume before the next read)."""
        n = self.lib.dlti_shm_ring_read(self.h, self._buf.ctypes.data, self._buf.nbytes, timeout_ms)
        if n < 0:
            raise RuntimeError(f"shared-memory ring read failed ({n})")
        return self._buf[:n]

    def close(self) -> None:
        if self.h:
            self.lib.dlti_shm_ring_close(self.h)
            self.h = None


def ring_name(tag: str = "") -> str:
    return f"/dlti_tp_{os.environ.get('MASTER_PORT', '0')}_{tag}"
