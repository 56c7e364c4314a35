"""h5min: the pure-Python HDF5 subset reader/writer behind load_model (reference predict.py:24).

h5py / libhdf5 are not installable in this image, so the reader is validated against (1) files produced by
the writer, including multi-level group B-trees and global-heap strings, and (2) hand-checked raw structure
offsets of the HDF5 spec (superblock v0, symbol-table entry, object header v1)."""
import os
import struct

import numpy as np
import pytest

import cnn_itmo_b200  # noqa: F401
from cnn_itmo_b200 import h5min


def _make(tmp_path, leaf_k=4, internal_k=16, n_groups=70):
    w = h5min.Writer(leaf_k=leaf_k, internal_k=internal_k)
    rng = np.random.default_rng(0)
    data = {}
    for i in range(n_groups):
        name = f'layer_{i}'
        a = rng.standard_normal((3, 3, 2, i % 5 + 1)).astype(np.float32)
        data[name] = a
        w.dataset(f'model_weights/{name}/{name}/kernel:0', a)
        w.group(f'model_weights/{name}').attrs['weight_names'] = np.array([f'{name}/kernel:0'.encode()])
    w.group('model_weights').attrs['layer_names'] = np.array([n.encode() for n in data])
    w.root.attrs['model_config'] = h5min.VlenStr(b'{"x": 1}' * 5000)    # 40 kB vlen string in the global heap
    w.root.attrs['keras_version'] = h5min.VlenStr(b'2.2.4')
    w.root.attrs['scalar'] = np.float32(2.5)
    w.root.attrs['ints'] = np.arange(5, dtype=np.int64)
    p = str(tmp_path / 'f.h5')
    w.save(p)
    return p, data


@pytest.mark.parametrize('leaf_k,internal_k', [(4, 16), (1, 1), (2, 2)])
def test_roundtrip_groups_datasets_attributes(tmp_path, leaf_k, internal_k):
    p, data = _make(tmp_path, leaf_k, internal_k)
    f = h5min.File(p)
    assert f.keys() == ['model_weights']
    mw = f['model_weights']
    assert mw.keys() == sorted(data)                       # B-tree traversal finds every child, any depth
    assert [n.decode() for n in mw.attrs['layer_names']] == list(data)
    for name, a in data.items():
        d = mw[f'{name}/{name}/kernel:0']
        assert d.is_dataset and d.shape == a.shape and d.dtype == np.float32
        np.testing.assert_array_equal(d.read(), a)
        assert mw[name].attrs['weight_names'][0] == f'{name}/kernel:0'.encode()
    assert f.attrs['keras_version'] == b'2.2.4'
    assert f.attrs['model_config'] == b'{"x": 1}' * 5000
    assert f.attrs['scalar'] == np.float32(2.5)
    np.testing.assert_array_equal(f.attrs['ints'], np.arange(5))
    assert 'model_weights/layer_3' in f and 'nope' not in f
    with pytest.raises(KeyError):
        f['model_weights/missing']


def test_raw_structures_follow_the_hdf5_spec(tmp_path):
    p, _ = _make(tmp_path, n_groups=3)
    raw = open(p, 'rb').read()
    assert raw[:8] == b'\x89HDF\r\n\x1a\n'
    assert raw[8] == 0 and raw[13] == 8 and raw[14] == 8                  # superblock v0, 8-byte offsets/lengths
    leaf_k, internal_k = struct.unpack_from('<HH', raw, 16)
    assert (leaf_k, internal_k) == (4, 16)
    base, free, eof, drv = struct.unpack_from('<QQQQ', raw, 24)
    assert base == 0 and eof == len(raw) and free == drv == h5min.UNDEF
    name_off, ohdr, cache, _ = struct.unpack_from('<QQII', raw, 56)       # root symbol-table entry
    btree, heap = struct.unpack_from('<QQ', raw, 80)
    assert cache == 1 and raw[btree:btree + 4] == b'TREE' and raw[heap:heap + 4] == b'HEAP'
    ver, _, nmsg, refc, hsize = struct.unpack_from('<BBHII', raw, ohdr)   # object header v1 prefix
    assert ver == 1 and refc == 1 and nmsg >= 1 and hsize % 8 == 0
    mtype, msize = struct.unpack_from('<HH', raw, ohdr + 16)              # first message: symbol table
    assert mtype == h5min.MSG_SYMBOL_TABLE and msize == 16
    assert struct.unpack_from('<QQ', raw, ohdr + 24) == (btree, heap)
    assert b'GCOL' in raw and b'SNOD' in raw


def test_errors_are_oserror(tmp_path):
    with pytest.raises(OSError):
        h5min.File(str(tmp_path / 'missing.h5'))
    bad = tmp_path / 'bad.h5'
    bad.write_bytes(b'not hdf5 at all' * 100)
    with pytest.raises(OSError):
        h5min.File(str(bad))
    p, _ = _make(tmp_path, n_groups=3)
    raw = open(p, 'rb').read()
    trunc = tmp_path / 'trunc.h5'
    trunc.write_bytes(raw[:len(raw) // 2])
    with pytest.raises(OSError):
        f = h5min.File(str(trunc))
        for k in f['model_weights'].keys():
            f['model_weights'][k][k]['kernel:0'].read()


def test_empty_and_fixed_string_arrays(tmp_path):
    w = h5min.Writer()
    g = w.group('g')
    g.attrs['weight_names'] = np.zeros((0,), dtype=np.float64)             # what Keras writes for weightless layers
    g.attrs['names'] = np.array([b'a', b'bcd', b'ef'])
    w.dataset('g/u8', np.arange(12, dtype=np.uint8).reshape(3, 4))
    w.dataset('g/f64', np.linspace(0, 1, 7))
    p = str(tmp_path / 'e.h5')
    w.save(p)
    f = h5min.File(p)
    assert f['g'].attrs['weight_names'].shape == (0,)
    assert list(f['g'].attrs['names']) == [b'a', b'bcd', b'ef']
    np.testing.assert_array_equal(f['g/u8'].read(), np.arange(12, dtype=np.uint8).reshape(3, 4))
    np.testing.assert_array_equal(f['g/f64'].read(), np.linspace(0, 1, 7))


# ---- files h5min did NOT write: hand-assembled from the HDF5 specification (tests/golden/make_h5py_style.py) -------------
@pytest.mark.parametrize('leaf_k,internal_k,chunked_every,cont', [(1, 2, 3, True), (4, 16, 0, True), (2, 1, 1, False)])
def test_h5py_style_checkpoint_is_read_by_load_model_reader(tmp_path, leaf_k, internal_k, chunked_every, cont):
    """A Keras-2.2.4 / h5py-2.x style file of the whole U-Net built by an independent assembler: multi-level group
    B-trees in strcmp (not layer) order, attributes in object-header continuation blocks, fixed-length string scalars
    and arrays, a vlen string through the global heap, a zero-length float64 weight_names, contiguous and chunked
    (overhanging edge chunks) float32 datasets, messages to skip. Read back through the reader load_model uses."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location(
        'make_h5py_style', os.path.join(os.path.dirname(__file__), 'golden', 'make_h5py_style.py'))
    mk = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mk)
    from cnn_itmo_b200 import checkpoint, graph
    from oracle import unet_ref as R
    ws = R.synthetic_weights(31)
    data = mk.build_keras_file(ws, graph.keras_layers(), checkpoint.model_config(), checkpoint.training_config(),
                               leaf_k=leaf_k, internal_k=internal_k, chunked_every=chunked_every,
                               attrs_in_continuation=cont)
    p = tmp_path / 'h5py_style.hdf5'
    p.write_bytes(data)
    got, meta = checkpoint.read_checkpoint(str(p))
    assert len(got) == graph.NUM_LAYERS
    for l, a, b in zip(graph.LAYERS, got, ws):
        for key in b:
            assert np.array_equal(a[key], np.asarray(b[key], np.float32)), (l.name, key)
    assert meta['keras_version'] == '2.2.4' and meta['backend'] == 'tensorflow'
    assert meta['model_config']['config']['name'] == graph.MODEL_NAME
    assert meta['training_config']['loss'] == 'mse'
    assert abs(meta['bn_epsilon'] - graph.BN_EPS) < 1e-12
    # the raw group order is strcmp order, which differs from the model order the reader must follow
    f = h5min.File(str(p))
    raw_order = f['model_weights'].keys()
    assert raw_order == sorted(raw_order) and raw_order != [k.name for k in graph.keras_layers()]
    f.close()
