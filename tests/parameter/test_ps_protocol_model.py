"""Exhaustive check of the chunk protocol of the striped parameter server (csrc/ps_kernels.cu) on a small model.

The kernels implement, per chunk: writers that count themselves in (``started += 1``, the same atomic returns the gate),
add to the data, and count themselves out (``finished += 1``); readers that copy between two looks at the counters and
fall back to raising the gate when they keep losing. This test runs EVERY interleaving of the atomic steps of a few
writers and readers (sequentially consistent steps - what the fences and atomics of the kernels provide) and checks the
two promises of the design:

* a reader never accepts a copy in which one element has seen a push that the other has not (no half-applied push);
* every push lands exactly once, every thread terminates (a writer that steps back for a gate retries and gets in).
"""
import pytest

ELEMS = 2  # elements of the modelled chunk; a push adds 1 to each, one element per step


def _initial(n_writers, n_readers):
    shared = (0, 0, 0) + (0,) * ELEMS                       # started, finished, gate, data...
    writers = tuple(('W0',) for _ in range(n_writers))      # (pc,)
    readers = tuple(('R0', 0, 0, 0) + (0,) * ELEMS for _ in range(n_readers))  # (pc, f, s, attempts, copy...)
    return shared, writers, readers


def _writer_step(shared, w):
    """One atomic step of a writer; returns (shared', w') or None when the step is not enabled (spinning)."""
    started, finished, gate, *data = shared
    pc = w[0]
    if pc == 'W0':                                         # started += 1, the same atomic reads the gate
        nxt = 'WB' if gate else 'W1'
        return (started + 1, finished, gate, *data), (nxt,)
    if pc.startswith('W') and pc[1:].isdigit() and 1 <= int(pc[1:]) <= ELEMS:   # red.add on element k
        k = int(pc[1:]) - 1
        data[k] += 1
        nxt = 'WF' if k == ELEMS - 1 else 'W%d' % (k + 2)
        return (started, finished, gate, *data), (nxt,)
    if pc == 'WF':                                         # fence, finished += 1
        return (started, finished + 1, gate, *data), ('done',)
    if pc == 'WB':                                         # stepped back: count out without having written ...
        return (started, finished + 1, gate, *data), ('WW',)
    if pc == 'WW':                                         # ... wait for the gate to come down, then try again
        return (shared, ('W0',)) if gate == 0 else None
    return None


def _reader_step(shared, r, retries, torn):
    started, finished, gate, *data = shared
    pc, f, s, attempts, *copy = r
    if pc == 'R0':                                         # finished first ...
        return shared, ('R1', finished, s, attempts, *copy)
    if pc == 'R1':                                         # ... then started: equal = nobody inside
        if started != f:
            return shared, ('R0', f, s, attempts, *copy)   # a writer is inside: look again
        return shared, ('C0', f, started, attempts, *copy)
    if pc.startswith('C'):                                 # copy element k
        k = int(pc[1:])
        copy[k] = data[k]
        nxt = 'R4' if k == ELEMS - 1 else 'C%d' % (k + 1)
        return shared, (nxt, f, s, attempts, *copy)
    if pc == 'R4':                                         # the version re-read
        if started == s:
            if len(set(copy)) != 1:
                torn.append(('optimistic', tuple(copy)))
            return shared, ('done', f, s, attempts, *copy)
        attempts += 1
        return shared, ('G0' if attempts >= retries else 'R0', f, s, attempts, *copy)
    if pc == 'G0':                                         # out of attempts: raise the gate
        return (started, finished, gate + 1, *data), ('G1', f, s, attempts, *copy)
    if pc == 'G1':
        return shared, ('G2', finished, s, attempts, *copy)
    if pc == 'G2':
        if started != f:
            return shared, ('G1', f, s, attempts, *copy)
        return shared, ('D0', f, started, attempts, *copy)
    if pc.startswith('D'):                                 # copy behind the gate
        k = int(pc[1:])
        copy[k] = data[k]
        nxt = 'G5' if k == ELEMS - 1 else 'D%d' % (k + 1)
        return shared, (nxt, f, s, attempts, *copy)
    if pc == 'G5':                                         # lower the gate
        if len(set(copy)) != 1:
            torn.append(('gated', tuple(copy)))
        return (started, finished, gate - 1, *data), ('done', f, s, attempts, *copy)
    return None


def _explore(n_writers, n_readers, retries):
    torn, finals, seen = [], set(), set()
    stack = [_initial(n_writers, n_readers)]
    while stack:
        state = stack.pop()
        if state in seen:
            continue
        seen.add(state)
        shared, writers, readers = state
        moved = False
        for i, w in enumerate(writers):
            if w[0] == 'done':
                continue
            out = _writer_step(shared, w)
            if out is not None:
                moved = True
                stack.append((tuple(out[0]), writers[:i] + (out[1],) + writers[i + 1:], readers))
        for i, r in enumerate(readers):
            if r[0] == 'done':
                continue
            out = _reader_step(shared, r, retries, torn)
            if out is not None:
                moved = True
                stack.append((tuple(out[0]), writers, readers[:i] + (tuple(out[1]),) + readers[i + 1:]))
        if not moved:
            finals.add(state)
    return torn, finals, len(seen)


@pytest.mark.parametrize('n_writers,n_readers,retries', [(1, 1, 1), (2, 1, 1), (2, 1, 2), (3, 1, 1), (2, 2, 1)])
def test_every_interleaving_keeps_the_promises(n_writers, n_readers, retries):
    torn, finals, states = _explore(n_writers, n_readers, retries)
    assert not torn, torn[:3]
    assert finals and states > 50
    for shared, writers, readers in finals:   # nothing is stuck: all threads done, every push landed once, gate down
        started, finished, gate, *data = shared
        assert all(w[0] == 'done' for w in writers) and all(r[0] == 'done' for r in readers)
        assert data == [n_writers] * ELEMS and started == finished and gate == 0


def test_the_model_catches_a_broken_protocol(monkeypatch):
    """The checker is not vacuous: a reader that skips the version re-read accepts torn copies."""
    import tests.parameter.test_ps_protocol_model as model
    original = model._reader_step

    def careless(shared, r, retries, torn):
        if r[0] == 'R4':   # accept whatever was copied
            if len(set(r[4:])) != 1:
                torn.append(('careless', tuple(r[4:])))
            return shared, ('done',) + tuple(r[1:])
        return original(shared, r, retries, torn)

    monkeypatch.setattr(model, '_reader_step', careless)
    torn, _, _ = model._explore(2, 1, 1)
    assert torn
