"""``SharedIO``: same-host job payloads through POSIX shared memory instead of the socket (reference:
``veles/txzmq/sharedio.py:45-105``: a memory-mapped region sized to the last payload + 5 %, re-created when a payload
outgrows it; the master and a slave on the same machine exchange only the segment's name over ZeroMQ)."""
from __future__ import annotations

from multiprocessing import shared_memory


class SharedIO(object):
    """A named shared-memory segment carrying one pickled payload between processes on one host."""

    def __init__(self, name=None, size=0, create=False):
        self.name = name
        self.size = size
        self._shm = shared_memory.SharedMemory(name=name, create=create, size=size if create else 0)
        if not create:
            self.size = self._shm.size

    @staticmethod
    def put(name, payload):
        try:
            old = shared_memory.SharedMemory(name=name)
            if old.size >= len(payload) + 8:
                shm = old
            else:
                old.close()
                old.unlink()
                shm = shared_memory.SharedMemory(name=name, create=True, size=int(len(payload) * 1.05) + 8)
        except FileNotFoundError:
            shm = shared_memory.SharedMemory(name=name, create=True, size=int(len(payload) * 1.05) + 8)
        shm.buf[:8] = len(payload).to_bytes(8, "little")
        shm.buf[8:8 + len(payload)] = payload
        shm.close()
        return name

    @staticmethod
    def get(name, unlink=False):
        shm = shared_memory.SharedMemory(name=name)
        try:
            n = int.from_bytes(bytes(shm.buf[:8]), "little")
            return bytes(shm.buf[8:8 + n])
        finally:
            shm.close()
            if unlink:
                try:
                    shm.unlink()
                except FileNotFoundError:
                    pass

    @staticmethod
    def remove(name):
        try:
            shm = shared_memory.SharedMemory(name=name)
            shm.close()
            shm.unlink()
        except FileNotFoundError:
            pass
