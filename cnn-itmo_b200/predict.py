"""Load -> predict -> save flow of the reference's predict.py (lines 24, 45-72), on the B200 path.

    python -m cnn_itmo_b200.predict                       # same defaults as the reference script
    python -m cnn_itmo_b200.predict --model ckpt.hdf5 --input-dir in --output-dir out --batch 16

Per file, as in the reference: PIL decode (predict.py:55) -> u8/255 (:59) -> model.predict (:62) ->
``(pred * 255).astype('uint8')`` truncation (:64) -> PNG (:68-70). The /255 and the truncation run fused
inside the first and last kernels (``Model.predict_u8``); ``--faithful`` goes through the float
``model.predict`` call instead, one image per call exactly like the reference loop.

Difference kept on purpose: the reference builds the output name with ``img.split('\\\\')[-1]`` (:70), which
only strips the directory on Windows; here the basename is used on every OS.
"""
from __future__ import annotations

import argparse
import glob
import os
import sys
from typing import List

import numpy as np

from . import checkpoint, graph
from .model import Model, load_model


def list_inputs(input_dir: str) -> List[str]:
    return sorted(glob.glob(os.path.join(input_dir, '*.png')))  # predict.py:46


def load_png(path: str) -> np.ndarray:
    from PIL import Image
    with Image.open(path) as im:  # predict.py:55
        a = np.asarray(im)
    if a.ndim == 2:
        a = np.stack([a] * 3, axis=-1)
    if a.shape[2] == 4:
        a = a[:, :, :3]
    if a.shape[:2] != (graph.IMG, graph.IMG):
        raise ValueError(f'{path}: {a.shape[1]}x{a.shape[0]} image; the model takes {graph.IMG}x{graph.IMG} '
                         '(use cnn_itmo_b200.frames for larger frames)')
    return np.ascontiguousarray(a, dtype=np.uint8)


def run(model: Model, files: List[str], output_dir: str, batch: int = 1, faithful: bool = False,
        prefix: str = '', verbose: bool = True) -> List[str]:
    from PIL import Image
    os.makedirs(output_dir, exist_ok=True)
    written: List[str] = []
    for i0 in range(0, len(files), batch):
        chunk = files[i0:i0 + batch]
        if verbose:
            print('Grabbing ', len(chunk), ' input files')  # predict.py:52
        imgs = np.stack([load_png(p) for p in chunk])
        if faithful:
            outs = []
            for im in imgs:
                X = np.asarray([np.true_divide(im.astype(float), 255)])   # predict.py:59-61
                pred = model.predict(X)                                    # predict.py:62
                outs.append((pred * 255)[0].astype('uint8'))               # predict.py:64
            out = np.stack(outs)
        else:
            out = model.predict_u8(imgs)
        for p, o in zip(chunk, out):
            dst = os.path.join(output_dir, prefix + os.path.basename(p))
            Image.fromarray(o).save(dst)                                   # predict.py:68-70
            written.append(dst)
    return written


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split('\n')[0])
    ap.add_argument('--model', default=None, help=f'Keras HDF5 checkpoint (default: {checkpoint.REFERENCE_CHECKPOINT})')
    ap.add_argument('--input-dir', default='images_to_predict/input')
    ap.add_argument('--output-dir', default='images_to_predict/output')
    ap.add_argument('--batch', type=int, default=1, help='images per predict call (reference: 1)')
    ap.add_argument('--precision', default=None, choices=['bf16', 'fp16', 'tf32', 'fp32'])
    ap.add_argument('--device', type=int, default=0)
    ap.add_argument('--faithful', action='store_true', help='float predict() per image like predict.py:59-64')
    a = ap.parse_args(argv)
    path = a.model or checkpoint.find_reference_checkpoint()
    if path is None:
        print(f'error: {checkpoint.REFERENCE_CHECKPOINT} not found (set --model or $ITMO_CKPT)', file=sys.stderr)
        return 2
    model = load_model(path, precision=a.precision, device=a.device, max_batch=max(1, a.batch))
    files = list_inputs(a.input_dir)
    run(model, files, a.output_dir, batch=max(1, a.batch), faithful=a.faithful)
    return 0


if __name__ == '__main__':
    sys.exit(main())
