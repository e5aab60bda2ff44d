"""Property tests (hypothesis) of the pure-Python building blocks: whatever the shapes, what goes in comes out."""
import socket
import threading

import numpy as np
import torch
from hypothesis import given, settings
from hypothesis import strategies as st
from hypothesis.extra import numpy as hnp

from elephas_b200.models import Dense, Sequential
from elephas_b200.parallel.flat import FlatLayout
from elephas_b200.utils import functional_utils as fu
from elephas_b200.utils.sockets import receive, send

widths = st.lists(st.integers(min_value=1, max_value=70), min_size=2, max_size=5)


def _chain(ws):
    return Sequential([Dense(ws[1], input_dim=ws[0])] + [Dense(w) for w in ws[2:]])


@settings(max_examples=30, deadline=None)
@given(widths, st.integers(0, 2 ** 31 - 1))
def test_flat_layout_round_trips_keras_arrays_and_model_tensors(ws, seed):
    """pack_arrays / unpack_arrays (Keras layout, the list the reference moves around) and pack_model / unpack_to_model
    (the model's own tensors) are inverse to each other for any widths - padded leading dimensions included - and the
    padding stays zero."""
    model = _chain(ws)
    layout = FlatLayout(model)
    rng = np.random.default_rng(seed)
    arrays = [rng.normal(size=w.shape).astype(np.float32) for w in model.get_weights()]
    flat = layout.pack_arrays(arrays)
    assert flat.numel() == layout.numel and layout.numel % 64 == 0
    back = layout.unpack_arrays(flat)
    assert all(np.array_equal(a, b) for a, b in zip(arrays, back))
    assert float(flat.abs().sum()) == float(sum(np.abs(a).astype(np.float64).sum() for a in arrays)) or \
        np.isclose(float(flat.double().abs().sum()), sum(np.abs(a).astype(np.float64).sum() for a in arrays), rtol=1e-6)
    model.set_weights(arrays)
    again = layout.pack_model(model)
    assert torch.equal(again, flat)
    other = _chain(ws)
    layout.unpack_to_model(flat, other)
    assert all(np.array_equal(a, b) for a, b in zip(other.get_weights(), arrays))


@settings(max_examples=30, deadline=None)
@given(st.lists(hnp.arrays(np.float32, hnp.array_shapes(max_dims=3, max_side=6), elements=st.floats(-100, 100, width=32)),
                min_size=1, max_size=5))
def test_parameter_list_arithmetic(arrays):
    """functional_utils on lists of arrays: (a + b) - b == a, a - a == neutral, divide_by scales every array."""
    doubled = fu.add_params(arrays, arrays)
    assert all(np.array_equal(d, 2 * a) for d, a in zip(doubled, arrays))
    assert all(np.array_equal(s, a) for s, a in zip(fu.subtract_params(doubled, arrays), arrays))
    zero = fu.get_neutral(arrays)
    assert all(z.shape == a.shape and not z.any() for z, a in zip(zero, arrays))
    assert all(np.allclose(h, a / 4.0) for h, a in zip(fu.divide_by(arrays, 4), arrays))


@settings(max_examples=20, deadline=None)
@given(st.lists(hnp.arrays(np.float32, hnp.array_shapes(max_dims=2, max_side=40), elements=st.floats(-1, 1, width=32)),
                max_size=4), st.binary(max_size=3000))
def test_socket_frames_of_any_size(arrays, blob):
    """The 20-digit length header + pickle framing of the socket parameter server (elephas/utils/sockets.py:45-71) for
    payloads from empty to tens of kilobytes, sent while the other side reads."""
    a, b = socket.socketpair()
    payload = {'delta': arrays, 'blob': blob}
    t = threading.Thread(target=send, args=(a, payload))
    t.start()
    got = receive(b)
    t.join()
    a.close()
    b.close()
    assert got['blob'] == blob and len(got['delta']) == len(arrays)
    assert all(np.array_equal(x, y) and x.dtype == y.dtype for x, y in zip(got['delta'], arrays))
