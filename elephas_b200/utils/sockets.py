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


def _receive_all(sock, num_bytes: int) -> bytes:
    """Read exactly ``num_bytes`` from ``sock``; raises ``ConnectionError`` on EOF."""
    chunks = []
    remaining = num_bytes
    while remaining > 0:
        data = sock.recv(min(remaining, 1 << 20))
        if not data:
            raise ConnectionError("socket closed with %d bytes outstanding" % remaining)
        chunks.append(data)
        remaining -= len(data)
    return b''.join(chunks)


def receive(sock, num_bytes: int = HEADER_BYTES):
    """Receive one length-prefixed pickle frame."""
    length = int(_receive_all(sock, num_bytes).decode())
    return pickle.loads(_receive_all(sock, length))


def send(sock, data, num_bytes: int = HEADER_BYTES):
    """Send ``data`` as one length-prefixed pickle frame."""
    payload = pickle.dumps(data, -1)
    sock.sendall(str(len(payload)).zfill(num_bytes).encode())
    sock.sendall(payload)
