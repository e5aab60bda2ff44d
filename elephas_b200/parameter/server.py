"""Parameter servers: hold the master weights, serve reads, apply ``w <- w - delta``.

Reference: ``elephas/parameter/server.py:17-233``. Behavioural contract kept:

* ``mode='asynchronous'`` takes a reader/writer lock around reads and updates, ``'hogwild'`` takes
  none (``server.py:109-131,205-218``);
* HTTP routes ``/`` (liveness text ``Elephas``), ``/parameters`` (GET, pickled weights),
  ``/update`` (POST, pickled delta); socket opcodes ``g`` / ``u`` with length-prefixed pickles.

Differences, all deliberate: the HTTP server is a stdlib
``ThreadingHTTPServer`` running in a thread of the driver (Flask and a forked process in the
reference), so the final weights do not need a last HTTP pull; the socket server takes the write
lock *before* reading the weights (the reference reads first, ``server.py:204-206``, and can lose
updates), leaves its connection loop on EOF instead of spinning (``server.py:220-226``) and
acknowledges each update with one byte. ``DeviceServer`` (``'nvlink'``) keeps the master weights in
GPU memory; see ``elephas_b200/parallel/device_ps.py``.
"""
import abc
import pickle
import socket
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

import numpy as np

from ..utils.functional_utils import subtract_params
from ..utils.notebook_utils import is_running_in_notebook
from ..utils.rwlock import RWLock as Lock
from ..utils.sockets import determine_master, receive, send


class BaseParameterServer(abc.ABC):
    """``model`` is the dictionary produced by ``model_to_dict`` (architecture JSON + weights)."""

    def __init__(self, model, port: int, mode: str, **kwargs):
        self.port = port
        self.mode = mode
        self.custom_objects = kwargs.get('custom_objects')
        self._model_dict = model
        self.weights = [np.array(w, dtype=np.float32, copy=True) for w in model['weights']]
        self.lock = Lock()
        self._master_network = None

    @property
    def master_network(self):
        """A model carrying the current master weights (built on first use)."""
        from ..utils.serialization import dict_to_model
        if self._master_network is None:
            self._master_network = dict_to_model(self._model_dict, self.custom_objects)
        self._master_network.set_weights(self.get_weights())
        return self._master_network

    @property
    def locked(self):
        return self.mode == 'asynchronous'

    # thread-safe primitives shared by both transports
    def get_weights(self):
        if self.locked:
            with self.lock.reading():
                return [w.copy() for w in self.weights]
        return [w.copy() for w in self.weights]

    def set_weights(self, weights):
        if self.locked:
            with self.lock.writing():
                self.weights = [np.array(w, dtype=np.float32, copy=True) for w in weights]
        else:
            self.weights = [np.array(w, dtype=np.float32, copy=True) for w in weights]

    def apply_delta(self, delta):
        if self.locked:
            with self.lock.writing():
                self.weights = subtract_params(self.weights, delta)
        else:
            # hogwild: in-place, unsynchronised - concurrent pushes may interleave per array
            for w, d in zip(self.weights, delta):
                w -= d

    def _pickled_weights(self):
        if self.locked:
            with self.lock.reading():
                return pickle.dumps(self.weights, -1)
        return pickle.dumps(self.weights, -1)

    @abc.abstractmethod
    def start(self):
        """Start serving."""

    @abc.abstractmethod
    def stop(self):
        """Stop serving and release the port."""


class HttpServer(BaseParameterServer):
    def __init__(self, model, port: int, mode: str, **kwargs):
        super().__init__(model, port, mode, **kwargs)
        self.master_url = None
        notebook = is_running_in_notebook()
        self.debug = False if notebook else kwargs.get('debug', False)
        self.threaded = False if notebook else kwargs.get('threaded', True)
        self.use_reloader = False
        self._httpd = None
        self._thread = None

    def _handler_class(self):
        server = self

        class Handler(BaseHTTPRequestHandler):
            protocol_version = 'HTTP/1.0'

            def _reply(self, body, content_type='application/elephas'):
                self.send_response(200)
                self.send_header('Content-Type', content_type)
                self.send_header('Content-Length', str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def do_GET(self):
                if self.path == '/':
                    self._reply(b'Elephas', 'text/plain')
                elif self.path == '/parameters':
                    self._reply(server._pickled_weights())
                else:
                    self.send_error(404)

            def do_POST(self):
                if self.path != '/update':
                    self.send_error(404)
                    return
                length = int(self.headers.get('Content-Length', 0))
                delta = pickle.loads(self.rfile.read(length))
                server.apply_delta(delta)
                self._reply(b'Update done', 'text/plain')

            def log_message(self, fmt, *args):
                if server.debug:
                    BaseHTTPRequestHandler.log_message(self, fmt, *args)

        return Handler

    def start(self):
        if self._httpd is not None:
            self.stop()
        host = determine_master(self.port).split(':')[0]
        ThreadingHTTPServer.allow_reuse_address = True
        self._httpd = ThreadingHTTPServer((host, self.port), self._handler_class())
        self._httpd.daemon_threads = True
        self._thread = threading.Thread(target=self._httpd.serve_forever, kwargs={'poll_interval': 0.05},
                                        name='elephas-http-ps', daemon=True)
        self._thread.start()
        self.master_url = determine_master(self.port)

    def start_flask_service(self):
        """Blocking form of :meth:`start` under the reference's name (``elephas/parameter/server.py:91``, where it is
        the target of the server process): serves ``/``, ``/parameters`` and ``/update`` in the calling thread until
        :meth:`stop` is called from another one. The service is the stdlib HTTP server here, not Flask (D11)."""
        if self._httpd is not None:
            self.stop()
        host = determine_master(self.port).split(':')[0]
        ThreadingHTTPServer.allow_reuse_address = True
        self._httpd = ThreadingHTTPServer((host, self.port), self._handler_class())
        self._httpd.daemon_threads = True
        self.master_url = determine_master(self.port)
        self._httpd.serve_forever(poll_interval=0.05)

    def stop(self):
        if self._httpd is None:
            return
        self._httpd.shutdown()
        self._httpd.server_close()
        if self._thread is not None:
            self._thread.join()
        self._httpd = None
        self._thread = None


class SocketServer(BaseParameterServer):
    def __init__(self, model, port: int, mode: str, **kwargs):
        super().__init__(model, port, mode, **kwargs)
        self.socket = None
        self.runs = False
        self.connections = []
        self.thread = None

    def start(self):
        if self.thread is not None:
            self.stop()
        sock = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        sock.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        host = determine_master(port=self.port).split(':')[0]
        sock.bind((host, self.port))
        sock.listen(64)
        sock.settimeout(0.1)
        self.socket = sock
        self.runs = True
        self.thread = threading.Thread(target=self.run, name='elephas-socket-ps', daemon=True)
        self.thread.start()

    def stop(self):
        self.runs = False
        if self.thread is not None:
            self.thread.join()
            self.thread = None
        for t in self.connections:
            t.join()
        self.connections = []
        if self.socket is not None:
            self.socket.close()
            self.socket = None

    def start_server(self):
        """Blocking form of :meth:`start` under the reference's name (``elephas/parameter/server.py:173``, the target
        of its server thread): binds and runs the accept loop in the calling thread until :meth:`stop_server`."""
        sock = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        sock.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        host = determine_master(port=self.port).split(':')[0]
        sock.bind((host, self.port))
        sock.listen(64)
        sock.settimeout(0.1)
        self.socket = sock
        self.runs = True
        self.run()

    def stop_server(self):
        """Reference name of :meth:`stop` (``elephas/parameter/server.py:184``)."""
        self.stop()

    def update_parameters(self, conn, ack=False):
        data = receive(conn)
        self.apply_delta(data['delta'])
        if ack:  # opcode b'U' only: the reference's b'u' frame has no reply (elephas/parameter/server.py:201-210)
            try:
                conn.sendall(b'k')
            except OSError:
                pass

    def get_parameters(self, conn):
        send(conn, self.get_weights())

    OPCODE_TIMEOUT, PAYLOAD_TIMEOUT = 5.0, 120.0

    def action_listener(self, conn):
        """One connection: opcode ``g`` (send the weights), ``u`` (apply a delta, no reply: the reference's frame) or
        ``U`` (apply a delta, then acknowledge with one byte). The short timeout only paces the wait for the next
        opcode so that ``stop`` is noticed; a payload may take as long as it takes, and a connection that breaks in the
        middle of one is closed without touching the weights."""
        with conn:
            while self.runs:
                conn.settimeout(self.OPCODE_TIMEOUT)
                try:
                    op = conn.recv(1)
                except socket.timeout:
                    continue
                except OSError:
                    return
                if not op:
                    return  # peer closed the connection
                conn.settimeout(self.PAYLOAD_TIMEOUT)
                try:
                    if op in (b'u', b'U'):
                        self.update_parameters(conn, ack=op == b'U')
                    elif op == b'g':
                        self.get_parameters(conn)
                except (OSError, EOFError, ValueError, KeyError, pickle.UnpicklingError):
                    return  # half a frame: drop the connection, the master weights are untouched

    def run(self):
        while self.runs:
            try:
                conn, _ = self.socket.accept()
            except socket.timeout:
                continue
            except OSError:
                break
            t = threading.Thread(target=self.action_listener, args=(conn,), daemon=True)
            t.start()
            self.connections.append(t)


class DeviceServer(BaseParameterServer):
    """Master weights resident in GPU (peer) memory, one stripe per rank; ``parameter_server_mode='nvlink'``.

    Under ``torchrun`` this server is *collective*: every rank constructs it, and ``start`` / ``stop`` are called by all
    ranks together (each contributes the stripe it owns, rank 0's weights become the master). The reference's counterpart
    lives in the driver process only (``elephas/spark_model.py:75-79,196-197,232-233``)."""
    collective = True

    def __init__(self, model, port: int, mode: str, **kwargs):
        super().__init__(model, port, mode, **kwargs)
        self._store = None
        self.fused_pull = bool(kwargs.get('fused_pull', False))  # kept for callers of the first design; unused

    def start(self):
        from ..parallel.device_ps import create_store
        from ..parallel.runtime import get_runtime
        rt = get_runtime()
        if rt.distributed and rt.device.type != 'cuda':
            raise RuntimeError("parameter_server_mode='nvlink' keeps the master weights in GPU peer memory; ranks that are "
                               "separate CPU processes have nothing to map - use 'socket' or 'http' there")
        self._store = create_store(self.port, self.weights, self.mode, model=self.master_network, rt=rt,
                                   device=rt.device if rt.distributed else None)

    def stop(self):
        from ..parallel.device_ps import destroy_store
        if self._store is not None:
            self.weights = self._store.pull_arrays()
            status = self._store.watchdog_status()
            if self._store.rt is not None:
                self._store.rt.barrier()  # nobody unmaps a stripe another rank may still be reading
            destroy_store(self.port)
            self._store = None
            if status:
                raise RuntimeError("device parameter server: a waiter of a chunk lock gave up after 10 s on this rank's "
                                   "stripe (status %d); the run is not trustworthy" % status)

    def get_weights(self):
        if self._store is not None:
            return self._store.pull_arrays()
        return super().get_weights()

    def set_weights(self, weights):
        super().set_weights(weights)
        if self._store is not None:
            self._store.load_arrays(self.weights)
