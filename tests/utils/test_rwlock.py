"""The reference leaves this file as a TODO (tests/utils/test_rwlock.py:1)."""
import threading
import time

from elephas.utils.rwlock import RWLock


def test_many_readers_one_writer():
    lock, state, errors = RWLock(), {'readers': 0, 'writer': False}, []

    def reader():
        for _ in range(200):
            lock.acquire_read()
            state['readers'] += 0  # presence only
            if state['writer']:
                errors.append('reader saw writer')
            lock.release()

    def writer():
        for _ in range(100):
            lock.acquire_write()
            state['writer'] = True
            time.sleep(0)
            state['writer'] = False
            lock.release()

    threads = [threading.Thread(target=reader) for _ in range(4)] + [threading.Thread(target=writer) for _ in range(2)]
    [t.start() for t in threads]
    [t.join() for t in threads]
    assert not errors


def test_writer_priority_blocks_new_readers():
    lock = RWLock()
    lock.acquire_read()
    order = []
    w = threading.Thread(target=lambda: (lock.acquire_write(), order.append('w'), lock.release()))
    w.start()
    time.sleep(0.05)  # the writer is now waiting
    r = threading.Thread(target=lambda: (lock.acquire_read(), order.append('r'), lock.release()))
    r.start()
    time.sleep(0.05)
    assert order == []  # the new reader queues behind the waiting writer
    lock.release()
    w.join(); r.join()
    assert order == ['w', 'r']


def test_exact_sum_under_write_lock():
    lock, box = RWLock(), [0]

    def bump():
        for _ in range(2000):
            with lock.writing():
                box[0] += 1

    ts = [threading.Thread(target=bump) for _ in range(4)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert box[0] == 8000
