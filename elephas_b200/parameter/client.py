"""Parameter-server clients.

Reference: ``elephas/parameter/client.py:13-91``. ``HttpClient`` (GET ``/parameters``, POST
``/update``, pickle bodies, ``Content-Type: application/elephas``) and ``SocketClient`` (opcode
``b'g'`` / ``b'u'`` + length-prefixed pickle frames) are wire-compatible with the reference's
servers (``SocketClient(ack=False)``; the default adds an acknowledged update opcode ``b'U'``). ``DeviceClient`` is the B200 addition (``parameter_server_mode='nvlink'``): pull and push
are kernels that read and update master weights resident in (peer) HBM - no host round trip.
"""
import abc
import os
import pickle
import socket
import time
import urllib.error
import urllib.request

import numpy as np

from ..utils.sockets import determine_master, receive, send


def _timeout():
    """Seconds a request may take before the worker gives up on the server (``ELEPHAS_B200_PS_TIMEOUT``, default 120):
    a server that died must fail the fit, not hang it."""
    return float(os.environ.get('ELEPHAS_B200_PS_TIMEOUT', '120'))


def _refused(exc):
    reason = getattr(exc, 'reason', exc)
    return isinstance(reason, ConnectionRefusedError)


def _when_listening(request, attempts=6, first_wait=0.05):
    """Run ``request()``; a REFUSED connection (the server thread has not bound its port yet, or is restarting) is tried
    again with doubling waits. Nothing has been sent in that case, so repeating is safe for pushes too; any other failure -
    a reset in the middle of a request, a timeout - is raised: the server may have applied the update already."""
    wait = first_wait
    for attempt in range(attempts):
        try:
            return request()
        except (ConnectionRefusedError, urllib.error.URLError) as exc:
            if not _refused(exc) or attempt == attempts - 1:
                raise
            time.sleep(wait)
            wait *= 2


class BaseParameterClient(abc.ABC):
    """A client can fetch the current master weights and send a weight delta."""
    client_type = 'base'

    @classmethod
    def get_client(cls, client_type: str, port: int = 4000):
        for klass in _all_subclasses(cls):
            if klass.client_type == client_type:
                return klass(port)
        raise ValueError("Parameter server mode has to be either `http`, `socket` or `nvlink`, "
                         "got {}".format(client_type))

    @abc.abstractmethod
    def update_parameters(self, delta: list):
        """Apply ``master -= delta`` on the server."""

    @abc.abstractmethod
    def get_parameters(self):
        """Fetch the master weights."""


def _all_subclasses(cls):
    out = []
    for sub in cls.__subclasses__():
        out.append(sub)
        out.extend(_all_subclasses(sub))
    return out


class HttpClient(BaseParameterClient):
    client_type = 'http'

    def __init__(self, port: int = 4000):
        self.master_url = determine_master(port=port)
        self.headers = {'Content-Type': 'application/elephas'}

    def _open(self, path, body=None):
        request = urllib.request.Request('http://%s/%s' % (self.master_url, path), data=body, headers=self.headers)

        def once():
            with urllib.request.urlopen(request, timeout=_timeout()) as response:
                return response.read()

        return _when_listening(once)

    def get_parameters(self):
        return pickle.loads(self._open('parameters'))

    def update_parameters(self, delta: list):
        return self._open('update', pickle.dumps(_to_numpy_list(delta), -1))


class SocketClient(BaseParameterClient):
    """Opens one TCP connection per request, like the reference (``client.py:78,86``)."""
    client_type = 'socket'

    def __init__(self, port: int = 4000, ack: bool = True):
        """``ack=True`` (default): updates use opcode ``U`` and wait for this package's server to confirm that the delta
        has been applied, so a pull that follows a push always sees it. ``ack=False``: the reference's un-acknowledged
        ``u`` frame - use it against a reference ``SocketServer`` (``elephas/parameter/server.py:220-226``), which
        never replies."""
        self.port = port
        self.ack = ack

    def _connect(self):
        host = determine_master(port=self.port).split(':')[0]
        def once():
            sock = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
            sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            sock.settimeout(_timeout())
            try:
                sock.connect((host, self.port))
            except BaseException:
                sock.close()
                raise
            return sock

        return _when_listening(once)

    def get_parameters(self):
        with self._connect() as sock:
            sock.sendall(b'g')
            return receive(sock)

    def update_parameters(self, delta: list):
        with self._connect() as sock:
            sock.sendall(b'U' if self.ack else b'u')
            send(sock, {'delta': _to_numpy_list(delta)})
            if self.ack and sock.recv(1) != b'k':
                raise ConnectionError("socket parameter server closed the connection before acknowledging an update")


def _to_numpy_list(arrays):
    out = []
    for a in arrays:
        if hasattr(a, 'detach'):
            a = a.detach().cpu().numpy()
        out.append(np.asarray(a))
    return out


class DeviceClient(BaseParameterClient):
    """Client of the device-resident parameter server (see ``elephas_b200/parallel/device_ps.py``).

    ``get_parameters`` / ``update_parameters`` keep the list-of-arrays contract of the other clients;
    GPU workers use ``pull_into`` / ``push_delta_from`` instead, which move flat device buffers with
    the ``ps_pull`` / ``ps_push`` kernels and never touch the host.
    """
    client_type = 'nvlink'

    def __init__(self, port: int = 4000):
        self.port = port

    def _store(self):
        from ..parallel.device_ps import get_store
        return get_store(self.port)

    def get_parameters(self):
        return self._store().pull_arrays()

    def update_parameters(self, delta: list):
        self._store().push_arrays(delta)
