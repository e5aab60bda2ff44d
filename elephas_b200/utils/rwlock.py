"""Writer-priority reader/writer lock.

Reference: ``elephas/utils/rwlock.py:1-67`` (same public methods: ``acquire_read``,
``acquire_write``, ``release``). Any number of readers may hold the lock at once, or exactly one
writer. A waiting writer blocks *new* readers so a stream of pulls cannot starve pushes. This host
lock guards the HTTP / socket parameter servers; the NVLink parameter server implements the same
policy on the device per chunk: writers count themselves in and out, readers validate their copy against the
counters and can raise a gate when they keep losing (``csrc/ps_kernels.cu``).
"""
import threading
from contextlib import contextmanager


class RWLock:
    def __init__(self):
        self._cond = threading.Condition(threading.Lock())
        self._active_readers = 0
        self._writer_active = False
        self._writers_waiting = 0

    def acquire_read(self):
        with self._cond:
            while self._writer_active or self._writers_waiting:
                self._cond.wait()
            self._active_readers += 1

    def acquire_write(self):
        with self._cond:
            self._writers_waiting += 1
            try:
                while self._writer_active or self._active_readers:
                    self._cond.wait()
            finally:
                self._writers_waiting -= 1
            self._writer_active = True

    def release(self):
        """Release the lock, whichever kind the caller holds."""
        with self._cond:
            if self._writer_active:
                self._writer_active = False
            elif self._active_readers:
                self._active_readers -= 1
            else:
                raise RuntimeError("release() called on an RWLock that is not held")
            self._cond.notify_all()

    @contextmanager
    def reading(self):
        self.acquire_read()
        try:
            yield self
        finally:
            self.release()

    @contextmanager
    def writing(self):
        self.acquire_write()
        try:
            yield self
        finally:
            self.release()
