"""Socket helpers of the TCP parameter server.

Reference: ``elephas/utils/sockets.py:6-71``. The wire format is kept byte-compatible: a frame is a
20-character, zero-padded ASCII decimal length followed by a pickle (highest protocol) of the
payload, so a reference ``SocketClient`` can talk to this server and vice versa.
"""
import os
import pickle
from socket import gethostbyname, gethostname

HEADER_BYTES = 20


def determine_master(port: int = 4000) -> str:
    """``host:port`` of the parameter server. ``SPARK_LOCAL_IP`` wins over hostname resolution
    (reference: ``sockets.py:18-21``); unresolvable hostnames fall back to loopback."""
    host = os.environ.get('SPARK_LOCAL_IP')
    if not host:
        try:
            host = gethostbyname(gethostname())
        except OSError:
            host = '127.0.0.1'
    return "%s:%d" % (host, port)


def _receive_all(socket, num_bytes: int) -> bytes:  # noqa: A002 (parameter named as in the reference)
    """Read exactly ``num_bytes`` from ``socket``; raises ``ConnectionError`` on EOF."""
    chunks = []
    remaining = num_bytes
    while remaining > 0:
        data = socket.recv(min(remaining, 1 << 20))
        if not data:
            raise ConnectionError("socket closed with %d bytes outstanding" % remaining)
        chunks.append(data)
        remaining -= len(data)
    return b''.join(chunks)


def receive(socket, num_bytes: int = HEADER_BYTES):
    """Receive one length-prefixed pickle frame."""
    length = int(_receive_all(socket, num_bytes).decode())
    return pickle.loads(_receive_all(socket, length))


def send(socket, data, num_bytes: int = HEADER_BYTES):
    """Send ``data`` as one length-prefixed pickle frame."""
    payload = pickle.dumps(data, -1)
    socket.sendall(str(len(payload)).zfill(num_bytes).encode())
    socket.sendall(payload)
