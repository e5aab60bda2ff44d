"""A small, dependency-free HDF5 subset: enough to write and read Keras-style checkpoints.

Why it exists: the reference saves models with ``keras.Model.save`` + ``h5py``
(``elephas/spark_model.py:115-125``, ``elephas/ml_model.py:61-70,172-186``); ``h5py`` is not part of
the B200 image and cannot be installed there. When ``h5py`` *is* importable it is used instead
(see ``open_backend``), so files interoperate either way.

Written format (HDF5 File Format Specification v3, all structures the HDF5 library >= 1.8 reads):
superblock version 2, version-2 object headers (``OHDR``, Jenkins lookup3 checksums), compact
"new style" groups (Link Info + Group Info + one Link message per child), contiguous little-endian
datasets (f4 / f8 / i4 / i8 / u1), version-3 attribute messages holding scalars, numeric arrays and
fixed-length byte strings (scalar or 1-D arrays) - the only attribute kinds Keras' HDF5 format uses.
An attribute larger than one object-header message can hold (64 KiB) is split into ``name__partK``
pieces that the reader reassembles.

Read support: everything this module writes, plus the classic layout ``h5py`` produces by default
(superblock 0/1, version-1 object headers with continuation blocks, symbol-table groups with
B-tree v1 / local heap, contiguous or compact datasets, fixed- and variable-length string
attributes through the global heap) - i.e. uncompressed Keras ``.h5`` files written elsewhere.
"""
import os
import struct
import threading

import numpy as np

SIGNATURE = b'\x89HDF\r\n\x1a\n'
UNDEF = 0xFFFFFFFFFFFFFFFF
_M = 0xFFFFFFFF
_MAX_ATTR_PAYLOAD = 60000


# ------------------------------------------------------------------ checksum
def _rot(x, k):
    return ((x << k) | (x >> (32 - k))) & _M


def lookup3(data: bytes, initval: int = 0) -> int:
    """Bob Jenkins' lookup3 ``hashlittle`` (the metadata checksum of HDF5 version-2 structures)."""
    length = len(data)
    a = b = c = (0xdeadbeef + length + initval) & _M
    i = 0
    while length > 12:
        a = (a + int.from_bytes(data[i:i + 4], 'little')) & _M
        b = (b + int.from_bytes(data[i + 4:i + 8], 'little')) & _M
        c = (c + int.from_bytes(data[i + 8:i + 12], 'little')) & _M
        a = (a - c) & _M; a ^= _rot(c, 4); c = (c + b) & _M
        b = (b - a) & _M; b ^= _rot(a, 6); a = (a + c) & _M
        c = (c - b) & _M; c ^= _rot(b, 8); b = (b + a) & _M
        a = (a - c) & _M; a ^= _rot(c, 16); c = (c + b) & _M
        b = (b - a) & _M; b ^= _rot(a, 19); a = (a + c) & _M
        c = (c - b) & _M; c ^= _rot(b, 4); b = (b + a) & _M
        i += 12
        length -= 12
    if length == 0:
        return c
    tail = data[i:] + b'\x00' * (12 - length)
    a = (a + int.from_bytes(tail[0:4], 'little')) & _M
    b = (b + int.from_bytes(tail[4:8], 'little')) & _M
    c = (c + int.from_bytes(tail[8:12], 'little')) & _M
    c ^= b; c = (c - _rot(b, 14)) & _M
    a ^= c; a = (a - _rot(c, 11)) & _M
    b ^= a; b = (b - _rot(a, 25)) & _M
    c ^= b; c = (c - _rot(b, 16)) & _M
    a ^= c; a = (a - _rot(c, 4)) & _M
    b ^= a; b = (b - _rot(a, 14)) & _M
    c ^= b; c = (c - _rot(b, 24)) & _M
    return c


# ---------------------------------------------------------------- data model
class Dataset:
    def __init__(self, data, attrs=None):
        self.data = np.asarray(data)
        self.attrs = dict(attrs or {})

    @property
    def shape(self):
        return self.data.shape

    def __getitem__(self, key):
        return self.data[key]

    def __array__(self, dtype=None, copy=None):
        return self.data.astype(dtype) if dtype is not None else self.data


class Group:
    def __init__(self, attrs=None):
        self.attrs = dict(attrs or {})
        self.children = {}

    def create_group(self, name):
        node = self
        for part in name.strip('/').split('/'):
            nxt = node.children.get(part)
            if nxt is None:
                nxt = node.children[part] = Group()
            node = nxt
        return node

    require_group = create_group

    def create_dataset(self, name, data, attrs=None):
        parts = name.strip('/').split('/')
        parent = self.create_group('/'.join(parts[:-1])) if len(parts) > 1 else self
        ds = parent.children[parts[-1]] = Dataset(data, attrs)
        return ds

    def __getitem__(self, name):
        node = self
        for part in name.strip('/').split('/'):
            node = node.children[part]
        return node

    def __contains__(self, name):
        try:
            self[name]
            return True
        except KeyError:
            return False

    def keys(self):
        return self.children.keys()

    def items(self):
        return self.children.items()


# -------------------------------------------------------------------- writer
def _dtype_message(dt: np.dtype) -> bytes:
    dt = np.dtype(dt)
    if dt.kind == 'f' and dt.itemsize in (4, 8):
        if dt.itemsize == 4:
            return bytes([0x11, 0x20, 0x1F, 0x00]) + struct.pack('<I', 4) + struct.pack('<HHBBBBI', 0, 32, 23, 8, 0, 23, 127)
        return bytes([0x11, 0x20, 0x3F, 0x00]) + struct.pack('<I', 8) + struct.pack('<HHBBBBI', 0, 64, 52, 11, 0, 52, 1023)
    if dt.kind in 'iu' and dt.itemsize in (1, 2, 4, 8):
        signed = 0x08 if dt.kind == 'i' else 0x00
        return bytes([0x10, signed, 0x00, 0x00]) + struct.pack('<I', dt.itemsize) + struct.pack('<HH', 0, dt.itemsize * 8)
    if dt.kind == 'S':
        # fixed-length string, null padded (what h5py writes for numpy bytes), ASCII character set
        return bytes([0x13, 0x01, 0x00, 0x00]) + struct.pack('<I', max(dt.itemsize, 1))
    raise TypeError("unsupported dtype for the built-in HDF5 writer: %r" % (dt,))


def _dataspace_message(shape) -> bytes:
    if shape == ():
        return bytes([2, 0, 0, 0])  # version 2, rank 0, no flags, scalar
    return bytes([2, len(shape), 0, 1]) + b''.join(struct.pack('<Q', int(d)) for d in shape)


def _normalise_attr(value):
    """-> little-endian NumPy array of a supported dtype (bytes / str become fixed-length strings)."""
    if isinstance(value, str):
        value = value.encode('utf8')
    if isinstance(value, bytes):
        return np.array(value if value else b'\x00', dtype='S%d' % max(len(value), 1))
    arr = np.asarray(value)
    if arr.dtype.kind == 'U':
        arr = np.char.encode(arr, 'utf8')
    if arr.dtype.kind == 'O':
        arr = np.array([v.encode('utf8') if isinstance(v, str) else v for v in arr.ravel()]).reshape(arr.shape)
    if arr.dtype.kind == 'b':
        arr = arr.astype(np.uint8)
    if arr.dtype.kind == 'f' and arr.dtype.itemsize == 2:
        arr = arr.astype(np.float32)
    if arr.dtype.kind == 'S' and arr.dtype.itemsize == 0:
        arr = arr.astype('S1')
    return arr.astype(arr.dtype.newbyteorder('<')) if arr.dtype.byteorder == '>' else arr


def _attribute_message(name: str, arr: np.ndarray) -> bytes:
    dt = _dtype_message(arr.dtype)
    ds = _dataspace_message(arr.shape)
    nm = name.encode('utf8') + b'\x00'
    return struct.pack('<BBHHHB', 3, 0, len(nm), len(dt), len(ds), 1 if any(b > 127 for b in nm) else 0) \
        + nm + dt + ds + np.ascontiguousarray(arr).tobytes()


def _attr_messages(attrs):
    msgs = []
    for name, value in attrs.items():
        arr = _normalise_attr(value)
        if arr.nbytes > _MAX_ATTR_PAYLOAD and arr.shape == () and arr.dtype.kind == 'S':
            raw = arr.tobytes()
            parts = [raw[i:i + _MAX_ATTR_PAYLOAD] for i in range(0, len(raw), _MAX_ATTR_PAYLOAD)]
            msgs.append((0x0C, _attribute_message(name + '__parts', np.array(len(parts), dtype=np.int64))))
            for k, chunk in enumerate(parts):
                msgs.append((0x0C, _attribute_message('%s__part%d' % (name, k), np.array(chunk, dtype='S%d' % len(chunk)))))
            continue
        if arr.nbytes > _MAX_ATTR_PAYLOAD:
            raise ValueError("attribute %r is too large for the built-in HDF5 writer (%d bytes)" % (name, arr.nbytes))
        msgs.append((0x0C, _attribute_message(name, arr)))
    return msgs


def _object_header(messages) -> bytes:
    body = b''.join(struct.pack('<BHB', mtype, len(data), 0) + data for mtype, data in messages)
    head = b'OHDR' + bytes([2, 0x02]) + struct.pack('<I', len(body))
    blob = head + body
    return blob + struct.pack('<I', lookup3(blob))


class _Writer:
    def __init__(self):
        self.buf = bytearray(48)  # superblock placeholder

    def _align(self, n=8):
        pad = (-len(self.buf)) % n
        self.buf += b'\x00' * pad

    def _append(self, blob):
        self._align()
        addr = len(self.buf)
        self.buf += blob
        return addr

    def write_dataset(self, ds: Dataset) -> int:
        # (np.asarray(..., order='C'), not ascontiguousarray: the latter turns a scalar dataset into shape (1,))
        arr = _normalise_attr(ds.data) if ds.data.dtype.kind in 'SUO' else np.asarray(ds.data, order='C')
        if arr.dtype.byteorder == '>':
            arr = arr.astype(arr.dtype.newbyteorder('<'))
        if arr.dtype.kind == 'f' and arr.dtype.itemsize == 2:
            arr = arr.astype(np.float32)
        raw = arr.tobytes()
        data_addr = self._append(raw) if raw else UNDEF
        msgs = [(0x01, _dataspace_message(arr.shape)), (0x03, _dtype_message(arr.dtype)),
                (0x05, bytes([3, 0x09])),
                (0x08, bytes([3, 1]) + struct.pack('<QQ', data_addr, len(raw)))]
        msgs += _attr_messages(ds.attrs)
        return self._append(_object_header(msgs))

    def write_group(self, grp: Group) -> int:
        links = []
        for name, child in grp.children.items():
            addr = self.write_group(child) if isinstance(child, Group) else self.write_dataset(child)
            nm = name.encode('utf8')
            if len(nm) < 256:
                links.append((0x06, bytes([1, 0x00, len(nm)]) + nm + struct.pack('<Q', addr)))
            else:
                links.append((0x06, bytes([1, 0x01]) + struct.pack('<H', len(nm)) + nm + struct.pack('<Q', addr)))
        msgs = [(0x02, bytes([0, 0]) + struct.pack('<QQ', UNDEF, UNDEF)), (0x0A, bytes([0, 0]))]
        msgs += links
        msgs += _attr_messages(grp.attrs)
        return self._append(_object_header(msgs))

    def finish(self, root_addr) -> bytes:
        self._align()
        sb = SIGNATURE + bytes([2, 8, 8, 0]) + struct.pack('<QQQQ', 0, UNDEF, len(self.buf), root_addr)
        sb += struct.pack('<I', lookup3(sb))
        self.buf[0:48] = sb
        return bytes(self.buf)


def write_file(path, root: Group):
    w = _Writer()
    root_addr = w.write_group(root)
    with open(path, 'wb') as fh:
        fh.write(w.finish(root_addr))


# -------------------------------------------------------------------- reader
class _Reader:
    def __init__(self, data: bytes):
        self.d = data
        self._gheaps = {}
        self.O = 8  # size of offsets
        self.L = 8  # size of lengths

    def u(self, off, n):
        return int.from_bytes(self.d[off:off + n], 'little')

    # ---- superblock
    def root(self):
        base = self.d.find(SIGNATURE)
        if base != 0:
            for cand in (512, 1024, 2048):
                if self.d[cand:cand + 8] == SIGNATURE:
                    base = cand
                    break
        if base < 0:
            raise ValueError("not an HDF5 file")
        ver = self.d[base + 8]
        if ver in (2, 3):
            self.O, self.L = self.d[base + 9], self.d[base + 10]
            root_addr = self.u(base + 12 + 3 * self.O, self.O)
            return self.read_object(root_addr)
        if ver in (0, 1):
            self.O, self.L = self.d[base + 13], self.d[base + 14]
            p = base + 24 + (4 if ver == 1 else 0)
            p += 4 * self.O  # base, free-space, eof, driver-info addresses
            # root group symbol table entry: link name offset, object header address, ...
            root_addr = self.u(p + self.O, self.O)
            return self.read_object(root_addr)
        raise ValueError("unsupported HDF5 superblock version %d" % ver)

    # ---- object headers
    def _messages_v1(self, addr):
        nmsgs = self.u(addr + 2, 2)
        size = self.u(addr + 8, 4)
        blocks = [(addr + 16, size)]
        msgs = []
        while blocks and len(msgs) < nmsgs:
            p, remaining = blocks.pop(0)
            end = p + remaining
            while p + 8 <= end and len(msgs) < nmsgs:
                mtype, msize, mflags = self.u(p, 2), self.u(p + 2, 2), self.d[p + 4]
                body = p + 8
                if mtype == 0x10:  # continuation
                    blocks.append((self.u(body, self.O), self.u(body + self.O, self.L)))
                msgs.append((mtype, body, msize, mflags))
                p = body + msize
        return msgs

    def _messages_v2(self, addr):
        flags = self.d[addr + 5]
        p = addr + 6
        if flags & 0x20:
            p += 16
        if flags & 0x10:
            p += 4
        szlen = 1 << (flags & 3)
        chunk = self.u(p, szlen)
        p += szlen
        blocks = [(p, chunk)]
        msgs = []
        track = bool(flags & 0x04)
        while blocks:
            p, size = blocks.pop(0)
            end = p + size
            while p + 4 <= end:
                mtype, msize, mflags = self.d[p], self.u(p + 1, 2), self.d[p + 3]
                body = p + 4 + (2 if track else 0)
                if mtype == 0x10:
                    caddr, clen = self.u(body, self.O), self.u(body + self.O, self.L)
                    blocks.append((caddr + 4, clen - 8))  # skip 'OCHK' signature and trailing checksum
                if mtype != 0 or msize:
                    msgs.append((mtype, body, msize, mflags))
                p = body + msize
        return msgs

    def read_object(self, addr):
        if self.d[addr:addr + 4] == b'OHDR':
            msgs = self._messages_v2(addr)
        else:
            msgs = self._messages_v1(addr)
        attrs, links, dataspace, dtype, layout, symtab = {}, [], None, None, None, None
        for mtype, body, msize, mflags in msgs:
            if mtype == 0x0C:
                name, value = self._attribute(body)
                attrs[name] = value
            elif mtype == 0x06:
                links.append(self._link(body))
            elif mtype == 0x01:
                dataspace = self._dataspace(body)[0]
            elif mtype == 0x03:
                dtype = self._datatype(body)[0]
            elif mtype == 0x08:
                layout = body
            elif mtype == 0x11:
                symtab = (self.u(body, self.O), self.u(body + self.O, self.O))
        attrs = _join_parts(attrs)
        if layout is not None and dtype is not None:
            return Dataset(self._dataset_values(layout, dataspace, dtype), attrs)
        grp = Group(attrs)
        for name, child_addr in links:
            grp.children[name] = self.read_object(child_addr)
        if symtab is not None:
            for name, child_addr in self._symbol_table(*symtab):
                grp.children[name] = self.read_object(child_addr)
        return grp

    # ---- messages
    def _link(self, p):
        flags = self.d[p + 1]
        q = p + 2
        ltype = 0
        if flags & 0x08:
            ltype = self.d[q]
            q += 1
        if flags & 0x04:
            q += 8
        if flags & 0x10:
            q += 1
        nlen_size = 1 << (flags & 3)
        nlen = self.u(q, nlen_size)
        q += nlen_size
        name = self.d[q:q + nlen].decode('utf8')
        q += nlen
        if ltype != 0:
            raise ValueError("only hard links are supported")
        return name, self.u(q, self.O)

    def _dataspace(self, p):
        ver, rank, flags = self.d[p], self.d[p + 1], self.d[p + 2]
        q = p + (8 if ver == 1 else 4)
        dims = tuple(self.u(q + i * self.L, self.L) for i in range(rank))
        q += rank * self.L * (2 if flags & 1 else 1)
        if ver == 2 and self.d[p + 3] == 2:
            dims = (0,)
        return dims, q - p

    def _datatype(self, p):
        cls, ver = self.d[p] & 0x0F, self.d[p] >> 4
        bits0 = self.d[p + 1]
        size = self.u(p + 4, 4)
        if cls == 0:
            kind = 'i' if bits0 & 0x08 else 'u'
            return (np.dtype('%s%s%d' % ('>' if bits0 & 1 else '<', kind, size)), None), 8 + 4
        if cls == 1:
            return (np.dtype('%sf%d' % ('>' if bits0 & 1 else '<', size)), None), 8 + 12
        if cls == 3:
            return (np.dtype('S%d' % size), None), 8
        if cls == 9:  # variable length; base type follows
            base, blen = self._datatype(p + 8)
            is_string = (bits0 & 0x0F) == 1
            return (np.dtype('O'), ('vlen_str' if is_string else 'vlen', base[0])), 8 + blen
        raise ValueError("unsupported HDF5 datatype class %d" % cls)

    def _global_heap_object(self, heap_addr, index):
        heap = self._gheaps.get(heap_addr)
        if heap is None:
            heap = {}
            size = self.u(heap_addr + 8, self.L)
            p, end = heap_addr + 8 + self.L, heap_addr + size
            while p + 8 + self.L <= end:
                idx, osize = self.u(p, 2), self.u(p + 8, self.L)
                if idx == 0:
                    break
                heap[idx] = self.d[p + 8 + self.L:p + 8 + self.L + osize]
                p += 8 + self.L + ((osize + 7) // 8) * 8
            self._gheaps[heap_addr] = heap
        return heap.get(index, b'')

    def _decode(self, raw, shape, dtype):
        dt, extra = dtype
        count = int(np.prod(shape)) if shape else 1
        if extra is not None:  # variable-length: (length, heap address, index) triples
            step = 4 + self.O + 4
            out = []
            for i in range(count):
                rec = raw[i * step:(i + 1) * step]
                obj = self._global_heap_object(int.from_bytes(rec[4:4 + self.O], 'little'),
                                               int.from_bytes(rec[4 + self.O:], 'little'))
                out.append(obj.decode('utf8') if extra[0] == 'vlen_str' else np.frombuffer(obj, dtype=extra[1]))
            return out[0] if shape == () else np.array(out, dtype=object).reshape(shape)
        arr = np.frombuffer(raw[:count * dt.itemsize], dtype=dt).reshape(shape)
        if dt.kind == 'S':
            return bytes(arr[()]).rstrip(b'\x00') if shape == () else arr.copy()
        if dt.byteorder == '>':
            arr = arr.astype(dt.newbyteorder('<'))
        return arr[()] if shape == () else arr.copy()

    def _attribute(self, p):
        ver = self.d[p]
        if ver == 1:
            nlen, dlen, slen = self.u(p + 2, 2), self.u(p + 4, 2), self.u(p + 6, 2)
            q = p + 8
            pad = lambda n: (n + 7) // 8 * 8  # noqa: E731
            name = self.d[q:q + nlen].split(b'\x00')[0].decode('utf8')
            q += pad(nlen)
            dtype, _ = self._datatype(q)
            q += pad(dlen)
            shape, _ = self._dataspace(q)
            q += pad(slen)
        else:
            nlen, dlen, slen = self.u(p + 2, 2), self.u(p + 4, 2), self.u(p + 6, 2)
            q = p + 8 + (1 if ver == 3 else 0)
            name = self.d[q:q + nlen].split(b'\x00')[0].decode('utf8')
            q += nlen
            dtype, _ = self._datatype(q)
            q += dlen
            shape, _ = self._dataspace(q)
            q += slen
        dt, extra = dtype
        item = (4 + self.O + 4) if extra is not None else dt.itemsize
        count = int(np.prod(shape)) if shape else 1
        return name, self._decode(self.d[q:q + count * item], shape, dtype)

    def _dataset_values(self, layout_p, shape, dtype):
        ver, cls = self.d[layout_p], self.d[layout_p + 1]
        if ver != 3:
            raise ValueError("unsupported data layout message version %d" % ver)
        shape = shape if shape is not None else ()
        dt, extra = dtype
        item = (4 + self.O + 4) if extra is not None else dt.itemsize
        count = int(np.prod(shape)) if shape else 1
        if cls == 0:  # compact
            size = self.u(layout_p + 2, 2)
            raw = self.d[layout_p + 4:layout_p + 4 + size]
        elif cls == 1:  # contiguous
            addr = self.u(layout_p + 2, self.O)
            raw = b'' if addr == UNDEF else self.d[addr:addr + count * item]
        else:
            raise ValueError("chunked / compressed datasets are not supported by the built-in HDF5 reader")
        if not raw and count:
            return np.zeros(shape, dtype=dt if extra is None else object)
        return np.asarray(self._decode(raw, shape, dtype))

    # ---- classic groups
    def _symbol_table(self, btree_addr, heap_addr):
        heap_data = self.u(heap_addr + 8 + 2 * self.L, self.O)

        def name_at(off):
            end = self.d.index(b'\x00', heap_data + off)
            return self.d[heap_data + off:end].decode('utf8')

        out = []

        def walk(addr):
            if self.d[addr:addr + 4] == b'TREE':
                level, used = self.d[addr + 5], self.u(addr + 6, 2)
                p = addr + 8 + 2 * self.O
                for i in range(used):
                    child = self.u(p + self.L + i * (self.L + self.O), self.O)
                    walk(child)
                _ = level
            elif self.d[addr:addr + 4] == b'SNOD':
                n = self.u(addr + 6, 2)
                p = addr + 8
                for i in range(n):
                    e = p + i * (2 * self.O + 24)
                    out.append((name_at(self.u(e, self.O)), self.u(e + self.O, self.O)))

        walk(btree_addr)
        return out


def _join_parts(attrs):
    joined = dict(attrs)
    for key in [k for k in attrs if k.endswith('__parts')]:
        base = key[:-len('__parts')]
        n = int(attrs[key])
        joined[base] = b''.join(bytes(attrs['%s__part%d' % (base, i)]) for i in range(n))
        del joined[key]
        for i in range(n):
            del joined['%s__part%d' % (base, i)]
    return joined


def read_file(path) -> Group:
    with open(path, 'rb') as fh:
        data = fh.read()
    root = _Reader(data).root()
    if not isinstance(root, Group):
        raise ValueError("HDF5 root object is not a group")
    return root


# --------------------------------------------------------- h5py interoperability
def have_h5py() -> bool:
    try:
        import h5py  # noqa: F401
        return True
    except Exception:
        return False


def save_tree(path, root: Group):
    """Write ``root`` with h5py when available, otherwise with the built-in writer. The file appears under its name in one
    step (written next to it, then renamed): a reader, or a second writer of the same checkpoint path - the workers of one
    process all run the same ``ModelCheckpoint`` - never sees a half-written file."""
    path = str(path)
    tmp = "%s.tmp-%d-%d" % (path, os.getpid(), threading.get_ident())
    try:
        _save_tree(tmp, root)
        os.replace(tmp, path)
    finally:
        if os.path.exists(tmp):
            os.remove(tmp)


def _save_tree(path, root: Group):
    if have_h5py():  # pragma: no cover - h5py is absent from the B200 image
        import h5py

        def emit(h5g, grp):
            for k, v in grp.attrs.items():
                h5g.attrs[k] = v
            for name, child in grp.children.items():
                if isinstance(child, Group):
                    emit(h5g.create_group(name), child)
                else:
                    d = h5g.create_dataset(name, data=child.data)
                    for k, v in child.attrs.items():
                        d.attrs[k] = v

        with h5py.File(path, 'w') as f:
            emit(f, root)
        return
    write_file(path, root)


def load_tree(path) -> Group:
    """Read a file into the in-memory tree (h5py when available, else the built-in reader)."""
    if have_h5py():  # pragma: no cover
        import h5py

        def absorb(h5g):
            grp = Group({k: (v if not isinstance(v, str) else v.encode('utf8')) for k, v in h5g.attrs.items()})
            for name, child in h5g.items():
                if isinstance(child, h5py.Group):
                    grp.children[name] = absorb(child)
                else:
                    grp.children[name] = Dataset(child[()], dict(child.attrs.items()))
            return grp

        with h5py.File(path, 'r') as f:
            return absorb(f)
    return read_file(path)
