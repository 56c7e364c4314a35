"""Stand-ins for the three Keras callbacks main.py wires into fit_generator (main.py:69, 121-124, 132):

    csv_logger  = CSVLogger('log.csv', append=True, separator=';')                                   # main.py:121
    checkpoint  = ModelCheckpoint("saved7-model-{epoch:02d}-{val_acc:.2f}.hdf5", verbose=1,
                                  save_best_only=False, mode='max')                                   # main.py:123-124
    testmodelcb = LambdaCallback(on_epoch_end=makePrediction)                                         # main.py:69

Same constructor arguments, same files on disk: CSVLogger writes ``epoch;<sorted log keys>`` once (not when appending to
an existing non-empty file) and one row per epoch; ModelCheckpoint formats the path with ``epoch + 1`` and the epoch's
logs and calls ``model.save`` (full Keras HDF5 with /optimizer_weights). Host-side code only.
"""
from __future__ import annotations

import csv
import os
from typing import Any, Callable, Dict, List, Optional

import numpy as np


class Callback:
    """The hook surface Model.fit_generator drives (a subset of keras.callbacks.Callback)."""

    def __init__(self):
        self.model = None
        self.params: Dict[str, Any] = {}

    def set_model(self, model):
        self.model = model

    def set_params(self, params):
        self.params = dict(params)

    def on_train_begin(self, logs=None):
        pass

    def on_train_end(self, logs=None):
        pass

    def on_epoch_begin(self, epoch, logs=None):
        pass

    def on_epoch_end(self, epoch, logs=None):
        pass


class LambdaCallback(Callback):
    """keras.callbacks.LambdaCallback (main.py:69 passes on_epoch_end=makePrediction)."""

    def __init__(self, on_epoch_begin: Optional[Callable] = None, on_epoch_end: Optional[Callable] = None,
                 on_train_begin: Optional[Callable] = None, on_train_end: Optional[Callable] = None, **kwargs):
        super().__init__()
        if on_epoch_begin is not None:
            self.on_epoch_begin = on_epoch_begin
        if on_epoch_end is not None:
            self.on_epoch_end = on_epoch_end
        if on_train_begin is not None:
            self.on_train_begin = on_train_begin
        if on_train_end is not None:
            self.on_train_end = on_train_end


class CSVLogger(Callback):
    """keras.callbacks.CSVLogger(filename, separator=',', append=False): one row per epoch, columns ``epoch`` + the
    sorted keys of the first epoch's logs; with append=True an existing file is continued without a second header."""

    def __init__(self, filename: str, separator: str = ',', append: bool = False):
        super().__init__()
        self.filename, self.sep, self.append = filename, separator, append
        self.keys: Optional[List[str]] = None
        self.append_header = True
        self._file = None
        self._writer = None

    def on_train_begin(self, logs=None):
        if self.append:
            if os.path.exists(self.filename):
                with open(self.filename, 'r') as f:
                    self.append_header = not bool(len(f.readline()))
            mode = 'a'
        else:
            mode = 'w'
        self._file = open(self.filename, mode, newline='')

    def on_epoch_end(self, epoch, logs=None):
        logs = logs or {}
        if self._file is None:
            self.on_train_begin()

        def fmt(v):
            if isinstance(v, (np.ndarray, list, tuple)):
                return '"[%s]"' % ', '.join(map(str, v))
            return v
        if self.keys is None:
            self.keys = sorted(logs.keys())
        if self._writer is None:
            self._writer = csv.DictWriter(self._file, fieldnames=['epoch'] + self.keys, delimiter=self.sep)
            if self.append_header:
                self._writer.writeheader()
        row = {'epoch': epoch}
        row.update((k, fmt(logs.get(k, 'NA'))) for k in self.keys)
        self._writer.writerow(row)
        self._file.flush()

    def on_train_end(self, logs=None):
        if self._file is not None:
            self._file.close()
        self._file = self._writer = None


class ModelCheckpoint(Callback):
    """keras.callbacks.ModelCheckpoint: ``filepath.format(epoch=epoch + 1, **logs)``; with save_best_only=False (main.py:124)
    the model is saved after every `period` epochs."""

    def __init__(self, filepath: str, monitor: str = 'val_loss', verbose: int = 0, save_best_only: bool = False,
                 save_weights_only: bool = False, mode: str = 'auto', period: int = 1):
        super().__init__()
        self.filepath, self.monitor, self.verbose = filepath, monitor, verbose
        self.save_best_only, self.save_weights_only, self.period = save_best_only, save_weights_only, period
        self.epochs_since_last_save = 0
        if mode not in ('auto', 'min', 'max'):
            mode = 'auto'
        if mode == 'min' or (mode == 'auto' and not ('acc' in monitor or monitor.startswith('fmeasure'))):
            self.monitor_op, self.best = np.less, np.inf
        else:
            self.monitor_op, self.best = np.greater, -np.inf

    def _save(self, path):
        if self.save_weights_only:
            self.model.save_weights(path, overwrite=True)
        else:
            self.model.save(path, overwrite=True)

    def on_epoch_end(self, epoch, logs=None):
        logs = logs or {}
        self.epochs_since_last_save += 1
        if self.epochs_since_last_save < self.period:
            return
        self.epochs_since_last_save = 0
        path = self.filepath.format(epoch=epoch + 1, **logs)
        if self.save_best_only:
            current = logs.get(self.monitor)
            if current is None:
                print(f'Can save best model only with {self.monitor} available, skipping.')
                return
            if not self.monitor_op(current, self.best):
                if self.verbose > 0:
                    print('\nEpoch %05d: %s did not improve from %0.5f' % (epoch + 1, self.monitor, self.best))
                return
            if self.verbose > 0:
                print('\nEpoch %05d: %s improved from %0.5f to %0.5f, saving model to %s'
                      % (epoch + 1, self.monitor, self.best, current, path))
            self.best = current
        elif self.verbose > 0:
            print('\nEpoch %05d: saving model to %s' % (epoch + 1, path))
        self._save(path)
