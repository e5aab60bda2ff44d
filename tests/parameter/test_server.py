"""The reference leaves this file as a TODO (tests/parameter/test_server.py:1). Servers and clients are
tested directly here: protocol round trips, the w <- w - delta arithmetic against a NumPy oracle, and
exactness of concurrent pushes in the locked ('asynchronous') mode."""
import threading
from itertools import count

import numpy as np
import pytest

from elephas.parameter.factory import ClientServerFactory
from elephas.utils.serialization import model_to_dict
from elephas_b200.models import Dense, Sequential

_ports = count(5200)


def _tiny_model():
    m = Sequential()
    m.add(Dense(8, input_dim=4, activation='relu'))
    m.add(Dense(2))
    return m


@pytest.mark.parametrize('ps', ['http', 'socket', 'nvlink'])
@pytest.mark.parametrize('mode', ['asynchronous', 'hogwild'])
def test_pull_push_round_trip(ps, mode):
    port = next(_ports)
    model = _tiny_model()
    factory = ClientServerFactory.get_factory(ps)
    server = factory.create_server(model_to_dict(model), port, mode)
    client = factory.create_client(port)
    server.start()
    try:
        w0 = model.get_weights()
        pulled = client.get_parameters()
        assert all(np.allclose(a, b) for a, b in zip(pulled, w0))
        delta = [np.full_like(w, 0.25) for w in w0]
        client.update_parameters(delta)
        client.update_parameters(delta)
        after = client.get_parameters()
        assert all(np.allclose(a, b - 0.5) for a, b in zip(after, w0))
        assert all(np.allclose(a, b - 0.5) for a, b in zip(server.get_weights(), w0))
    finally:
        server.stop()


@pytest.mark.parametrize('ps', ['http', 'socket', 'nvlink'])
def test_locked_mode_loses_no_update(ps):
    port = next(_ports)
    model = _tiny_model()
    factory = ClientServerFactory.get_factory(ps)
    server = factory.create_server(model_to_dict(model), port, 'asynchronous')
    server.start()
    try:
        w0 = model.get_weights()
        n_threads, n_push = 4, 10

        def pusher():
            client = factory.create_client(port)
            for _ in range(n_push):
                client.update_parameters([np.ones_like(w) for w in w0])

        threads = [threading.Thread(target=pusher) for _ in range(n_threads)]
        [t.start() for t in threads]
        [t.join() for t in threads]
        final = factory.create_client(port).get_parameters()
        for a, b in zip(final, w0):
            assert np.allclose(a, b - n_threads * n_push, atol=1e-4)
    finally:
        server.stop()


def test_unknown_factory():
    with pytest.raises(ValueError):
        ClientServerFactory.get_factory('smoke-signals')


def test_http_home_route():
    import urllib.request
    port = next(_ports)
    server = ClientServerFactory.get_factory('http').create_server(model_to_dict(_tiny_model()), port, 'hogwild')
    server.start()
    try:
        assert urllib.request.urlopen('http://%s/' % server.master_url).read() == b'Elephas'
    finally:
        server.stop()


def test_device_store_on_cpu_keeps_the_list_of_arrays_contract():
    """On CPU tensors the host RWLock does the exclusion; push / pull keep the reference's list-of-arrays contract."""
    import numpy as np
    from elephas_b200.models import Dense, Sequential
    from elephas_b200.parallel.device_ps import create_store, destroy_store
    import torch
    model = Sequential([Dense(4, input_dim=6), Dense(2)])
    weights = model.get_weights()
    store = create_store(45991, weights, 'asynchronous', model=model, device=torch.device('cpu'))
    try:
        assert store.locked and not store.on_device
        store.configure(local_workers=3, distributed=False)  # accepted for callers of the first design: nothing to choose
        delta = [np.ones_like(w) for w in weights]
        store.push_arrays(delta)
        for got, w in zip(store.pull_arrays(), weights):
            assert np.allclose(got, w - 1.0)
        assert store.pushes == 1 and store.matrix_source(0) is None and store.watchdog_status() == 0
    finally:
        destroy_store(45991)


def test_stripe_plan_covers_the_flat_buffer_exactly_once():
    """Ownership plan of the striped store: whole Dense kernels per owner, the vector tail split evenly, every flat
    element in exactly one span, stripe-local ranges disjoint, chunk ranges inside one chunk each."""
    import numpy as np
    from elephas_b200.models import Dense, Sequential, zoo
    from elephas_b200.parallel.device_ps import MAX_CHUNKS, MIN_CHUNK, StripePlan
    from elephas_b200.parallel.flat import FlatLayout
    for model in (zoo.mnist_mlp(), zoo.otto_mlp(), Sequential([Dense(4, input_dim=6), Dense(2)]),
                  zoo.wide_mlp(input_dim=512, hidden=512, depth=8, out_dim=1)):
        layout = FlatLayout(model)
        for world in (1, 2, 3, 8):
            plan = StripePlan(layout, world)
            seen = np.zeros(layout.numel, dtype=np.int32)
            used = [np.zeros(plan.capacity, dtype=np.int32) for _ in range(world)]
            for flat_off, n, owner, owner_off in plan.spans:
                assert n > 0 and n % 64 == 0 and flat_off % 64 == 0 and owner_off % 64 == 0 and 0 <= owner < world
                seen[flat_off:flat_off + n] += 1
                used[owner][owner_off:owner_off + n] += 1
            assert (seen == 1).all()
            for r in range(world):
                assert used[r].max() <= 1 and used[r].sum() == plan.stripe_len[r]
            assert sum(plan.stripe_len) == layout.numel
            assert plan.chunk_len % 64 == 0 and plan.chunk_len >= MIN_CHUNK and 1 <= plan.chunks <= MAX_CHUNKS
            assert plan.chunks * plan.chunk_len >= max(plan.stripe_len)
            # every Dense kernel is one span with a home; with several ranks the big model's layers spread out
            mats = [s for s in layout.segments if s._is_matrix]
            assert all(plan.matrix_home(s.offset) is not None for s in mats)
            if world == 8 and len(mats) >= 8:
                assert len({plan.matrix_home(s.offset)[0] for s in mats}) == 8
            covered = np.zeros(layout.numel, dtype=np.int32)
            for lo, hi in plan.chunk_ranges_flat():
                covered[lo:hi] += 1
            assert (covered == 1).all()


def test_stripe_plan_invariants_hold_for_arbitrary_dense_chains():
    """Property: for any chain of Dense layers (odd widths included) and any world size the plan puts every flat element
    in exactly one span, keeps stripe-local ranges disjoint and sizes the chunks within the lock-record budget."""
    import numpy as np
    from hypothesis import given, settings
    from hypothesis import strategies as st
    from elephas_b200.models import Dense, Sequential
    from elephas_b200.parallel.device_ps import MAX_CHUNKS, MIN_CHUNK, MAX_SPANS, StripePlan
    from elephas_b200.parallel.flat import FlatLayout

    @settings(max_examples=40, deadline=None, derandomize=True)
    @given(st.lists(st.integers(min_value=1, max_value=300), min_size=2, max_size=7), st.integers(min_value=1, max_value=8))
    def check(widths, world):
        model = Sequential([Dense(widths[1], input_dim=widths[0])] + [Dense(w) for w in widths[2:]])
        layout = FlatLayout(model)
        plan = StripePlan(layout, world)
        assert 1 <= len(plan.spans) <= MAX_SPANS
        seen = np.zeros(layout.numel, dtype=np.int32)
        used = [np.zeros(plan.capacity, dtype=np.int32) for _ in range(world)]
        for flat_off, n, owner, owner_off in plan.spans:
            assert n > 0 and n % 64 == 0 and flat_off % 64 == 0 and owner_off % 64 == 0 and 0 <= owner < world
            seen[flat_off:flat_off + n] += 1
            used[owner][owner_off:owner_off + n] += 1
        assert (seen == 1).all()
        for r in range(world):
            assert used[r].max() <= 1 and used[r].sum() == plan.stripe_len[r]
        assert plan.chunk_len % 64 == 0 and plan.chunk_len >= MIN_CHUNK and 1 <= plan.chunks <= MAX_CHUNKS
        assert plan.chunks * plan.chunk_len >= max(plan.stripe_len)
        covered = np.zeros(layout.numel, dtype=np.int32)
        for lo, hi in plan.chunk_ranges_flat():
            covered[lo:hi] += 1
        assert (covered == 1).all()

    check()


def test_reference_method_names_of_the_servers_work_as_blocking_services():
    """``HttpServer.start_flask_service`` and ``SocketServer.start_server`` / ``stop_server`` (the reference's entry
    points, run there as process / thread targets) serve in the calling thread until stopped."""
    import threading
    import time
    import numpy as np
    from elephas_b200.models import Dense, Sequential
    from elephas_b200.parameter.client import HttpClient, SocketClient
    from elephas_b200.parameter.server import HttpServer, SocketServer
    from elephas_b200.utils.serialization import model_to_dict
    model = Sequential([Dense(3, input_dim=4), Dense(1)])
    for cls, client_cls, port, entry, stop in ((HttpServer, HttpClient, 45981, 'start_flask_service', 'stop'),
                                               (SocketServer, SocketClient, 45982, 'start_server', 'stop_server')):
        server = cls(model_to_dict(model), port, 'asynchronous')
        t = threading.Thread(target=getattr(server, entry), daemon=True)
        t.start()
        client = client_cls(port)
        for _ in range(100):
            try:
                before = client.get_parameters()
                break
            except Exception:
                time.sleep(0.05)
        else:
            raise AssertionError('server did not come up')
        client.update_parameters([np.ones_like(w) for w in before])
        after = client.get_parameters()
        assert all(np.allclose(a, b - 1.0) for a, b in zip(after, before))
        if hasattr(client, 'close'):
            client.close()
        getattr(server, stop)()
        t.join(timeout=10)
        assert not t.is_alive()


def test_socket_server_speaks_the_reference_frames_and_survives_half_a_frame():
    """Opcode ``u`` is the reference's un-acknowledged update (elephas/parameter/client.py:85-91): nothing comes back on
    the wire. Opcode ``U`` is acknowledged. A client that dies in the middle of a payload costs the server one
    connection, not the weights and not the listener."""
    import socket as _socket
    import time
    from elephas_b200.parameter.client import SocketClient
    from elephas_b200.utils.sockets import determine_master, send
    port = next(_ports)
    model = _tiny_model()
    factory = ClientServerFactory.get_factory('socket')
    server = factory.create_server(model_to_dict(model), port, 'asynchronous')
    server.start()
    try:
        w0 = model.get_weights()
        delta = [np.full_like(w, 0.5) for w in w0]
        host = determine_master(port=port).split(':')[0]
        # the reference frame: no reply, the connection stays usable for the next opcode
        with _socket.create_connection((host, port)) as sock:
            sock.sendall(b'u')
            send(sock, {'delta': delta})
            sock.settimeout(0.3)
            with pytest.raises(_socket.timeout):
                sock.recv(1)
        legacy = SocketClient(port, ack=False)
        legacy.update_parameters(delta)
        # half a frame, then the client disappears
        with _socket.create_connection((host, port)) as sock:
            sock.sendall(b'U' + b'00000000000000009999' + b'x' * 10)
        acked = SocketClient(port)  # default: acknowledged updates
        acked.update_parameters(delta)
        deadline = time.time() + 5.0
        while time.time() < deadline:
            got = acked.get_parameters()
            if all(np.allclose(a, b - 1.5) for a, b in zip(got, w0)):
                break
            time.sleep(0.05)
        assert all(np.allclose(a, b - 1.5) for a, b in zip(acked.get_parameters(), w0))
    finally:
        server.stop()


def test_clients_wait_for_a_server_that_is_not_listening_yet_and_give_up_on_a_dead_one(monkeypatch):
    """A refused connection is retried (the server thread may not have bound its port yet); a server that never comes up,
    or one that accepts and never answers, fails the request instead of hanging the worker."""
    import socket
    import threading
    import time

    from elephas_b200.parameter import client as C

    with socket.socket() as probe:
        probe.bind(('127.0.0.1', 0))
        port = probe.getsockname()[1]
    monkeypatch.setenv('SPARK_LOCAL_IP', '127.0.0.1')
    started = time.time()
    with pytest.raises(ConnectionRefusedError):
        C.SocketClient(port).get_parameters()
    assert time.time() - started >= 0.05 * (1 + 2 + 4 + 8 + 16) * 0.9      # it did wait between its attempts

    calls = []

    def flaky():
        calls.append(1)
        if len(calls) < 3:
            raise ConnectionRefusedError()
        return 'up'

    assert C._when_listening(flaky) == 'up' and len(calls) == 3
    with pytest.raises(ConnectionResetError):   # anything but "refused" is not repeated: the update may have been applied
        C._when_listening(lambda: (_ for _ in ()).throw(ConnectionResetError()))

    silent = socket.socket()
    silent.bind(('127.0.0.1', 0))
    silent.listen(1)
    monkeypatch.setenv('ELEPHAS_B200_PS_TIMEOUT', '0.3')
    holder = []
    t = threading.Thread(target=lambda: holder.append(silent.accept()), daemon=True)
    t.start()
    with pytest.raises((socket.timeout, TimeoutError)):
        C.SocketClient(silent.getsockname()[1]).get_parameters()
    t.join(2)
    for conn, _ in holder:
        conn.close()
    silent.close()
