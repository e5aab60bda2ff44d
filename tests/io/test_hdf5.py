import struct

import numpy as np
import pytest

from elephas_b200.io import hdf5
from elephas_b200.models import SGD, load_model, to_categorical, zoo


def test_lookup3_vectors():
    # self-test vectors from Bob Jenkins' lookup3.c
    assert hdf5.lookup3(b'') == 0xdeadbeef
    assert hdf5.lookup3(b'Four score and seven years ago') == 0x17770551
    assert hdf5.lookup3(b'Four score and seven years ago', 1) == 0xcd628161


def test_tree_round_trip(tmp_path):
    root = hdf5.Group({'text': b'hello', 'number': np.int64(7), 'vec': np.arange(4, dtype=np.float64),
                       'names': np.array([b'dense/kernel:0', b'dense/bias:0'])})
    g = root.create_group('a/b')
    g.attrs['empty'] = np.zeros((0,))
    g.create_dataset('w:0', np.arange(12, dtype=np.float32).reshape(3, 4), attrs={'unit': b'none'})
    root.create_dataset('i64', np.arange(5, dtype=np.int64))
    path = str(tmp_path / 't.h5')
    hdf5.write_file(path, root)
    back = hdf5.read_file(path)
    assert back.attrs['text'] == b'hello' and int(back.attrs['number']) == 7
    assert np.array_equal(back.attrs['vec'], np.arange(4.0))
    assert [bytes(n) for n in back.attrs['names']] == [b'dense/kernel:0', b'dense/bias:0']
    assert np.array_equal(back['a/b/w:0'].data, np.arange(12, dtype=np.float32).reshape(3, 4))
    assert back['a/b/w:0'].attrs['unit'] == b'none' and back['a/b'].attrs['empty'].shape == (0,)
    assert np.array_equal(back['i64'].data, np.arange(5))


def test_file_structure_is_spec_conformant(tmp_path):
    path = str(tmp_path / 's.h5')
    hdf5.write_file(path, hdf5.Group({'k': b'v'}))
    raw = open(path, 'rb').read()
    assert raw[:8] == b'\x89HDF\r\n\x1a\n' and raw[8] == 2 and raw[9] == 8 and raw[10] == 8
    eof, root_addr = struct.unpack('<QQ', raw[28:44])
    assert eof == len(raw) and raw[root_addr:root_addr + 4] == b'OHDR'
    assert struct.unpack('<I', raw[44:48])[0] == hdf5.lookup3(raw[:44])
    size = struct.unpack('<I', raw[root_addr + 6:root_addr + 10])[0]
    end = root_addr + 10 + size
    assert struct.unpack('<I', raw[end:end + 4])[0] == hdf5.lookup3(raw[root_addr:end])


def test_large_attribute_is_split_and_rejoined(tmp_path):
    blob = bytes(range(1, 251)) * 700  # 175 kB
    path = str(tmp_path / 'big.h5')
    hdf5.write_file(path, hdf5.Group({'model_config': blob}))
    assert hdf5.read_file(path).attrs['model_config'] == blob


def test_keras_layout_and_model_round_trip(tmp_path, mnist_data):
    x, y, _, _ = mnist_data
    m = zoo.mnist_mlp(dropout=0.0)
    m.compile(SGD(lr=0.1, momentum=0.9), 'categorical_crossentropy', ['acc'])
    m.fit(x[:200], y[:200], batch_size=50, epochs=1)
    path = str(tmp_path / 'm.h5')
    m.save(path)
    tree = hdf5.read_file(path)
    assert set(tree.attrs) >= {'keras_version', 'backend', 'model_config', 'training_config'}
    mw = tree['model_weights']
    names = [bytes(n).decode() for n in mw.attrs['layer_names']]
    assert names == [l.name for l in m.layers]
    dense = m.layers[1].name
    assert mw[dense + '/' + dense + '/kernel:0'].data.shape == (784, 128)  # Keras (in, out) layout
    assert 'optimizer_weights' in tree
    # optimizer slots are stored in the Keras layout too (momentum of a Dense kernel: (in, out)), element for element
    ow = tree['optimizer_weights']
    slot = ow['SGD/' + dense + '/kernel/momentum:0'].data
    assert slot.shape == (784, 128)
    assert np.array_equal(slot, m.optimizer._slots[(0, 'momentum')].detach().cpu().numpy().T)
    m2 = load_model(path)
    assert np.array_equal(m2.optimizer._slots[(0, 'momentum')].cpu().numpy(),
                          m.optimizer._slots[(0, 'momentum')].cpu().numpy())
    assert m2.to_json() == m.to_json()
    assert np.array_equal(m2.predict(x[:20]), m.predict(x[:20]))
    assert m2.optimizer.get_config() == m.optimizer.get_config() and m2.optimizer.iterations == 4
    assert m2.loss == 'categorical_crossentropy' and m2.metrics_names == ['loss', 'acc']
    # resume: momentum slots survived
    s1 = m.train_on_batch(x[:50], y[:50]); s2 = m2.train_on_batch(x[:50], y[:50])
    assert np.allclose(m.get_weights()[-2], m2.get_weights()[-2], atol=1e-6)


def test_not_hdf5(tmp_path):
    p = tmp_path / 'x.h5'
    p.write_bytes(b'nope' * 100)
    with pytest.raises(ValueError):
        hdf5.read_file(str(p))


def test_reads_a_classic_libhdf5_layout_built_from_the_spec():
    """The committed fixture is a version-0 superblock file with symbol-table groups (TREE / HEAP / SNOD) and version-1
    object headers, assembled byte by byte from the HDF5 specification by ``make_v0_fixture.py`` - NOT by this
    package's writer (which emits the version-2 layout). It is the layout libhdf5 / h5py / Keras write by default."""
    import os
    from tests.io import make_v0_fixture
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'fixtures', 'classic_v0.h5')
    with open(path, 'rb') as fh:
        blob = fh.read()
    assert blob == make_v0_fixture.build()          # the committed bytes are what the generator states
    assert blob[8] == 0 and b'TREE' in blob and b'SNOD' in blob and b'HEAP' in blob and b'OHDR' not in blob
    root = hdf5.read_file(path)
    assert sorted(root.keys()) == ['g', 'w']
    assert root.attrs['note'] == b'built by hand' and int(root.attrs['answer']) == 42
    w = np.asarray(root['w'])
    assert w.dtype == np.float32 and np.array_equal(w, (np.arange(6, dtype=np.float32) * 0.5).reshape(2, 3))
    b = root['g']['b']
    assert np.array_equal(np.asarray(b), np.array([7, -1, 0, 2 ** 20], dtype=np.int32))
    assert float(b.attrs['scale']) == 0.25


def test_random_trees_round_trip(tmp_path):
    """Property: groups nested at random depths with datasets of random dtypes / shapes (scalars, empty and odd-sized
    arrays) and attributes come back exactly as written."""
    from hypothesis import given, settings
    from hypothesis import strategies as st
    from hypothesis.extra import numpy as hnp

    names = st.text(alphabet='abcdefgh_:0123456789', min_size=1, max_size=12)
    dtypes = st.sampled_from([np.float32, np.float64, np.int32, np.int64])
    arrays = dtypes.flatmap(lambda dt: hnp.arrays(dt, hnp.array_shapes(min_dims=0, max_dims=3, min_side=0, max_side=5),
                                                   elements=st.integers(-1000, 1000) if np.issubdtype(dt, np.integer)
                                                   else st.floats(-1e6, 1e6, width=32)))
    counter = [0]

    @settings(max_examples=25, deadline=None, derandomize=True)
    @given(st.dictionaries(st.lists(names, min_size=1, max_size=3).map('/'.join), arrays, min_size=1, max_size=6),
           st.dictionaries(names, arrays, max_size=3))
    def check(datasets, attrs):
        # a path can be a dataset or a group, not both: drop paths that are a prefix of another one
        paths = [p for p in datasets if not any(o != p and o.startswith(p + '/') for o in datasets)]
        root = hdf5.Group(dict(attrs))
        for p in paths:
            parent, _, leaf = p.rpartition('/')
            (root.create_group(parent) if parent else root).create_dataset(leaf, datasets[p])
        counter[0] += 1
        path = str(tmp_path / ('r%d.h5' % counter[0]))
        hdf5.write_file(path, root)
        back = hdf5.read_file(path)
        for p in paths:
            got = back[p].data
            assert got.dtype == datasets[p].dtype and got.shape == datasets[p].shape and np.array_equal(got, datasets[p])
        for k, v in attrs.items():
            assert np.array_equal(np.asarray(back.attrs[k]), v) and np.asarray(back.attrs[k]).dtype == v.dtype

    check()


def test_concurrent_writers_of_one_path_leave_a_whole_file(tmp_path):
    """Several workers checkpointing to one name: every reader finds a complete file, no temporary files stay behind."""
    import threading
    path = str(tmp_path / 'shared.h5')
    trees = []
    for i in range(4):
        root = hdf5.Group()
        root.attrs['writer'] = np.int64(i)
        root.create_dataset('payload', np.full((256, 64), float(i), dtype=np.float32))
        trees.append(root)
    errors = []

    def writer(tree):
        try:
            for _ in range(10):
                hdf5.save_tree(path, tree)
                back = hdf5.load_tree(path)
                who = int(np.asarray(back.attrs['writer']))
                assert np.all(np.asarray(back['payload'].data) == float(who))
        except Exception as exc:  # noqa: BLE001 - collected for the main thread
            errors.append(exc)

    threads = [threading.Thread(target=writer, args=(t,)) for t in trees]
    [t.start() for t in threads]
    [t.join() for t in threads]
    assert not errors, errors[:1]
    assert sorted(p.name for p in tmp_path.iterdir()) == ['shared.h5']
