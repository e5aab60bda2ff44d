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


def test_device_store_exclusion_policy_on_cpu():
    """``configure`` picks how mode='asynchronous' excludes writers; on CPU tensors the host RWLock does the work and
    the GPU-only shortcuts stay off."""
    import numpy as np
    from elephas_b200.models import Dense, Sequential
    from elephas_b200.parallel.device_ps import create_store, destroy_store
    import torch
    model = Sequential([Dense(4, input_dim=6), Dense(2)])
    weights = model.get_weights()
    store = create_store(45991, weights, 'asynchronous', model=model, device=torch.device('cpu'))
    try:
        assert store._exclusion == 'ticket+stream'
        store.configure(local_workers=3, distributed=False)
        assert store._exclusion == 'stream' and not store.stream_exclusive and not store.ticket_only  # CPU store
        store.configure(local_workers=1, distributed=True)
        assert store._exclusion == 'ticket'
        store.configure(local_workers=2, distributed=True)
        assert store._exclusion == 'ticket+stream'
        delta = [np.ones_like(w) for w in weights]
        store.push_arrays(delta)
        for got, w in zip(store.pull_arrays(), weights):
            assert np.allclose(got, w - 1.0)
        assert store.mirror is None and store.pushes == 1
    finally:
        destroy_store(45991)


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
