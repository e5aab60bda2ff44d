"""The reference leaves this file as a TODO (tests/utils/test_sockets.py:1); the framing is tested here."""
import socket
import threading

import numpy as np

from elephas.utils.sockets import determine_master, receive, send


def test_determine_master_env(monkeypatch):
    monkeypatch.setenv('SPARK_LOCAL_IP', '10.1.2.3')
    assert determine_master(4321) == '10.1.2.3:4321'


def test_frame_round_trip():
    a, b = socket.socketpair()
    payload = {'delta': [np.arange(100000, dtype=np.float32), np.ones((3, 3))]}
    t = threading.Thread(target=send, args=(a, payload))
    t.start()
    got = receive(b)
    t.join()
    assert np.array_equal(got['delta'][0], payload['delta'][0]) and got['delta'][1].shape == (3, 3)
    a.close(); b.close()


def test_header_is_20_ascii_digits():
    a, b = socket.socketpair()
    send(a, [1, 2, 3])
    header = b.recv(20)
    assert len(header) == 20 and header.isdigit()
    a.close(); b.close()
