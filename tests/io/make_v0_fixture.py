"""Builds ``fixtures/classic_v0.h5`` byte by byte from the HDF5 file-format specification (version-0 superblock,
version-1 object headers, symbol-table groups: ``TREE`` / ``HEAP`` / ``SNOD``) - the layout libhdf5 writes by default
and therefore what a Keras ``model.save`` produces. Nothing here imports ``elephas_b200.io.hdf5``: the fixture is an
independent statement of the format, and ``test_hdf5.py`` checks that the package's reader agrees with it.

    python tests/io/make_v0_fixture.py      # rewrites the committed fixture (bytes are deterministic)

Content:  /            attrs: note = b"built by hand" (fixed string), answer = 42 (int64)
          /w           float32 [2, 3] = [[0, 1, 2], [3, 4, 5]] * 0.5
          /g/b         int32 [4] = [7, -1, 0, 2**20],  attr scale = 0.25 (float64)
"""
import os
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
O = L = 8


def pad8(b):
    return b + b'\x00' * (-len(b) % 8)


def message(mtype, body):
    body = pad8(body)
    return struct.pack('<HHB3x', mtype, len(body), 0) + body


def object_header(messages):
    blob = b''.join(messages)
    return struct.pack('<BxHII4x', 1, len(messages), 1, len(blob)) + blob


def dataspace(shape):
    return struct.pack('<BBB5x', 1, len(shape), 0) + b''.join(struct.pack('<Q', d) for d in shape)


def dtype_float(size):
    exp_size, mant = {4: (8, 23), 8: (11, 52)}[size]
    return (struct.pack('<BBBBI', 0x11, 0x20, size * 8 - 1, 0, size) +
            struct.pack('<HHBBBBI', 0, size * 8, mant, exp_size, 0, mant, (1 << (exp_size - 1)) - 1))


def dtype_int(size, signed=True):
    return struct.pack('<BBBBI', 0x10, 0x08 if signed else 0, 0, 0, size) + struct.pack('<HH', 0, size * 8)


def dtype_string(size):
    return struct.pack('<BBBBI', 0x13, 0, 0, 0, size)


def attribute(name, dt, shape, raw):
    nm = name.encode() + b'\x00'
    sp = dataspace(shape)
    return struct.pack('<BxHHH', 1, len(nm), len(dt), len(sp)) + pad8(nm) + pad8(dt) + pad8(sp) + raw


class File:
    def __init__(self):
        self.buf = bytearray(96)  # the superblock is filled in last

    def put(self, blob):
        self.buf += b'\x00' * (-len(self.buf) % 8)
        addr = len(self.buf)
        self.buf += blob
        return addr

    def dataset(self, arr, dt, attrs=()):
        data = self.put(arr.tobytes())
        layout = struct.pack('<BB', 3, 1) + struct.pack('<QQ', data, arr.nbytes)
        msgs = [message(0x0001, dataspace(arr.shape)), message(0x0003, dt), message(0x0008, layout)]
        msgs += [message(0x000C, a) for a in attrs]
        return self.put(object_header(msgs))

    def group(self, children, attrs=()):
        """``children``: name -> object header address. Returns (header address, b-tree address, heap address)."""
        names = sorted(children)
        heap_data, offsets = b'\x00' * 8, {}
        for n in names:
            offsets[n] = len(heap_data)
            heap_data += pad8(n.encode() + b'\x00')
        data_addr = self.put(heap_data)
        heap = self.put(b'HEAP' + struct.pack('<B3xQQQ', 0, len(heap_data), UNDEF, data_addr))
        entries = b''.join(struct.pack('<QQII16x', offsets[n], children[n], 0, 0) for n in names)
        snod = self.put(b'SNOD' + struct.pack('<BxH', 1, len(names)) + entries)
        tree = self.put(b'TREE' + struct.pack('<BBHQQ', 0, 0, 1, UNDEF, UNDEF) +
                        struct.pack('<QQQ', 0, snod, offsets[names[-1]]))
        msgs = [message(0x0011, struct.pack('<QQ', tree, heap))] + [message(0x000C, a) for a in attrs]
        return self.put(object_header(msgs)), tree, heap

    def finish(self, root, tree, heap):
        sb = b'\x89HDF\r\n\x1a\n' + struct.pack('<BBBBBBBxHHI', 0, 0, 0, 0, 0, O, L, 4, 16, 0)
        sb += struct.pack('<QQQQ', 0, UNDEF, len(self.buf), UNDEF)
        sb += struct.pack('<QQII', 0, root, 1, 0) + struct.pack('<QQ', tree, heap)
        assert len(sb) == 96
        self.buf[:96] = sb
        return bytes(self.buf)


def build():
    f = File()
    w = f.dataset((np.arange(6, dtype='<f4') * 0.5).reshape(2, 3), dtype_float(4))
    b = f.dataset(np.array([7, -1, 0, 2 ** 20], dtype='<i4'), dtype_int(4),
                  attrs=[attribute('scale', dtype_float(8), (), struct.pack('<d', 0.25))])
    g, _, _ = f.group({'b': b})
    note = b'built by hand'
    root, tree, heap = f.group({'g': g, 'w': w}, attrs=[attribute('note', dtype_string(len(note)), (), note),
                                                         attribute('answer', dtype_int(8), (), struct.pack('<q', 42))])
    return f.finish(root, tree, heap)


if __name__ == '__main__':
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'fixtures', 'classic_v0.h5')
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, 'wb') as fh:
        fh.write(build())
    print(path, os.path.getsize(path), 'bytes')
