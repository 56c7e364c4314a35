"""Host-side pieces around the hot path that need no GPU: the three Keras callbacks main.py wires into fit_generator
(main.py:69, 121-124, 132), the PNG codec of the predict.py loop (predict.py:54-58, 66-70) and the pipelined loop's
ordering, driven with a stand-in model."""
import csv
import os

import numpy as np
import pytest

from cnn_itmo_b200 import callbacks as C
from cnn_itmo_b200 import predict as P


class FakeModel:
    """Records what the callbacks ask of it."""

    def __init__(self):
        self.saved, self.saved_weights = [], []

    def save(self, path, overwrite=True):
        self.saved.append(path)
        open(path, 'wb').write(b'full')

    def save_weights(self, path, overwrite=True):
        self.saved_weights.append(path)
        open(path, 'wb').write(b'weights')


def _rows(path, sep):
    with open(path, newline='') as f:
        return list(csv.reader(f, delimiter=sep))


def test_csv_logger_header_separator_and_append(tmp_path):
    # main.py:121  CSVLogger('log.csv', append=True, separator=';')
    path = str(tmp_path / 'log.csv')
    cb = C.CSVLogger(path, append=True, separator=';')
    cb.on_train_begin()
    cb.on_epoch_end(0, {'loss': 0.5, 'acc': 0.25, 'val_loss': 0.75, 'val_acc': 0.125})
    cb.on_epoch_end(1, {'loss': 0.25, 'acc': 0.5, 'val_loss': 0.5, 'val_acc': 0.25})
    cb.on_train_end()
    rows = _rows(path, ';')
    assert rows[0] == ['epoch', 'acc', 'loss', 'val_acc', 'val_loss']        # 'epoch' + sorted log keys
    assert rows[1] == ['0', '0.25', '0.5', '0.125', '0.75'] and rows[2][0] == '1'
    # a second run appends to the same file WITHOUT a second header (append=True)
    cb = C.CSVLogger(path, append=True, separator=';')
    cb.on_train_begin()
    cb.on_epoch_end(0, {'loss': 0.1, 'acc': 0.9, 'val_loss': 0.2, 'val_acc': 0.8})
    cb.on_train_end()
    rows = _rows(path, ';')
    assert len(rows) == 4 and rows[3] == ['0', '0.9', '0.1', '0.8', '0.2']
    assert sum(r[0] == 'epoch' for r in rows) == 1
    # append=False truncates
    cb = C.CSVLogger(path)
    cb.on_train_begin()
    cb.on_epoch_end(0, {'loss': 1.0})
    cb.on_train_end()
    assert _rows(path, ',') == [['epoch', 'loss'], ['0', '1.0']]


def test_csv_logger_missing_key_and_array_value(tmp_path):
    path = str(tmp_path / 'l.csv')
    cb = C.CSVLogger(path)
    cb.on_train_begin()
    cb.on_epoch_end(0, {'loss': 1.0, 'lr': np.array([0.1, 0.2])})
    cb.on_epoch_end(1, {'loss': 2.0})                                         # a key of epoch 0 missing later -> 'NA'
    cb.on_train_end()
    rows = _rows(path, ',')
    assert rows[0] == ['epoch', 'loss', 'lr'] and rows[2] == ['1', '2.0', 'NA']
    assert '0.1' in rows[1][2] and '0.2' in rows[1][2]


def test_model_checkpoint_reference_filename_pattern(tmp_path):
    # main.py:123-124  ModelCheckpoint("saved7-model-{epoch:02d}-{val_acc:.2f}.hdf5", verbose=1, save_best_only=False, mode='max')
    m = FakeModel()
    cb = C.ModelCheckpoint(str(tmp_path / 'saved7-model-{epoch:02d}-{val_acc:.2f}.hdf5'), verbose=0,
                           save_best_only=False, mode='max')
    cb.set_model(m)
    cb.on_epoch_end(0, {'val_acc': 0.456, 'loss': 1.0})
    cb.on_epoch_end(1, {'val_acc': 0.5, 'loss': 0.9})
    assert [os.path.basename(p) for p in m.saved] == ['saved7-model-01-0.46.hdf5', 'saved7-model-02-0.50.hdf5']
    assert all(os.path.exists(p) for p in m.saved) and not m.saved_weights


def test_model_checkpoint_best_only_period_and_weights_only(tmp_path):
    m = FakeModel()
    cb = C.ModelCheckpoint(str(tmp_path / 'best-{epoch:02d}.h5'), monitor='val_loss', save_best_only=True,
                           save_weights_only=True)
    cb.set_model(m)
    for e, v in enumerate([0.5, 0.6, 0.4, 0.45]):
        cb.on_epoch_end(e, {'val_loss': v})
    assert [os.path.basename(p) for p in m.saved_weights] == ['best-01.h5', 'best-03.h5'] and not m.saved
    # monitor 'val_acc' in auto mode maximises
    m = FakeModel()
    cb = C.ModelCheckpoint(str(tmp_path / 'acc-{epoch:02d}.h5'), monitor='val_acc', save_best_only=True)
    cb.set_model(m)
    for e, v in enumerate([0.5, 0.4, 0.7]):
        cb.on_epoch_end(e, {'val_acc': v})
    assert [os.path.basename(p) for p in m.saved] == ['acc-01.h5', 'acc-03.h5']
    # period=2 saves every second epoch
    m = FakeModel()
    cb = C.ModelCheckpoint(str(tmp_path / 'p-{epoch:02d}.h5'), period=2)
    cb.set_model(m)
    for e in range(5):
        cb.on_epoch_end(e, {'val_loss': 1.0})
    assert [os.path.basename(p) for p in m.saved] == ['p-02.h5', 'p-04.h5']


def test_lambda_callback_hooks():
    seen = []
    cb = C.LambdaCallback(on_epoch_end=lambda epoch, logs: seen.append((epoch, dict(logs))))   # main.py:69
    cb.on_train_begin({})
    cb.on_epoch_begin(0, {})
    cb.on_epoch_end(0, {'loss': 1.0})
    cb.on_train_end({})
    assert seen == [(0, {'loss': 1.0})]


# ---- PNG codec and the pipelined loop ------------------------------------------------------------------------------
def test_decode_png_equals_load_png_on_the_reference_inputs(golden_dir):
    files = P.list_inputs(os.path.join(golden_dir, 'images_to_predict', 'input'))
    assert len(files) == 12
    buf = np.empty((512, 512, 3), np.uint8)
    for p in files:
        a = P.load_png(p)
        b = P.decode_png(p)
        np.testing.assert_array_equal(a, b)
        c = P.decode_png(p, out=buf)                                          # decode straight into a batch buffer
        assert c is buf
        np.testing.assert_array_equal(a, buf)


def test_encode_png_round_trip(tmp_path):
    rng = np.random.RandomState(3)
    img = rng.randint(0, 256, (512, 512, 3)).astype(np.uint8)
    for level in (None, 1, 6):
        p = str(tmp_path / f'o{level}.png')
        P.encode_png(img, p, level)
        np.testing.assert_array_equal(P.load_png(p), img)


class _HostModel:
    """predict_u8 stand-in: a fixed per-pixel map, so the pipelined loop can be checked against run() without a GPU."""

    def predict_u8(self, imgs, out=None):
        res = (255 - np.asarray(imgs)).astype(np.uint8)
        if out is None:
            return res
        out[...] = res
        return out


def test_run_pipelined_writes_the_same_files_as_run(tmp_path, golden_dir, monkeypatch):
    files = P.list_inputs(os.path.join(golden_dir, 'images_to_predict', 'input'))
    # pinned host memory needs a CUDA context: plain numpy buffers here
    from cnn_itmo_b200 import model as model_mod
    monkeypatch.setattr(model_mod, 'pinned_empty', lambda shape, dtype: np.empty(shape, dtype))
    a_dir, b_dir = str(tmp_path / 'a'), str(tmp_path / 'b')
    m = _HostModel()
    wa = P.run(m, files, a_dir, batch=5, verbose=False)
    stats = {}
    wb = P.run_pipelined(m, files, b_dir, batch=5, workers=3, stats=stats)    # ragged last chunk (12 = 5 + 5 + 2)
    assert [os.path.basename(p) for p in wa] == [os.path.basename(p) for p in wb]
    for pa, pb in zip(wa, wb):
        np.testing.assert_array_equal(P.load_png(pa), P.load_png(pb))
    assert stats['files'] == 12 and stats['files_per_s'] > 0 and stats['batch'] == 5
    # a broken input surfaces as the worker's exception, not as a silently skipped file
    bad = str(tmp_path / 'bad.png')
    open(bad, 'wb').write(b'not a png')
    with pytest.raises(Exception):
        P.run_pipelined(m, files[:2] + [bad], str(tmp_path / 'c'), batch=2, workers=2)
