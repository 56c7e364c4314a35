"""Load -> predict -> save flow of the reference's predict.py (lines 24, 45-72), on the B200 path.

    python -m cnn_itmo_b200.predict                       # same defaults as the reference script
    python -m cnn_itmo_b200.predict --model ckpt.hdf5 --input-dir in --output-dir out --batch 16

Per file, as in the reference: PIL decode (predict.py:55) -> u8/255 (:59) -> model.predict (:62) ->
``(pred * 255).astype('uint8')`` truncation (:64) -> PNG (:68-70). The /255 and the truncation run fused
inside the first and last kernels (``Model.predict_u8``); ``--faithful`` goes through the float
``model.predict`` call instead, one image per call exactly like the reference loop. With ``--batch N`` (N > 1) the
files go through ``run_pipelined``: a thread pool decodes / encodes PNGs into / out of pinned buffers while the GPU
works on the chunk in between.

Difference kept on purpose: the reference builds the output name with ``img.split('\\\\')[-1]`` (:70), which
only strips the directory on Windows; here the basename is used on every OS.
"""
from __future__ import annotations

import argparse
import glob
import os
import sys
from typing import Dict, List, Optional

import numpy as np

from . import checkpoint, graph
from .model import Model, load_model


def list_inputs(input_dir: str) -> List[str]:
    return sorted(glob.glob(os.path.join(input_dir, '*.png')))  # predict.py:46


try:  # OpenCV's PNG codec releases the GIL, so a thread pool scales; PIL's decoder holds it
    import cv2 as _cv2  # type: ignore
    _cv2.setNumThreads(1)
except Exception:  # noqa: BLE001
    _cv2 = None


def _check_image(a: np.ndarray, path: str) -> np.ndarray:
    if a.ndim == 2:
        a = np.stack([a] * 3, axis=-1)
    if a.shape[2] == 4:
        a = a[:, :, :3]
    if a.shape[:2] != (graph.IMG, graph.IMG):
        raise ValueError(f'{path}: {a.shape[1]}x{a.shape[0]} image; the model takes {graph.IMG}x{graph.IMG} '
                         '(use cnn_itmo_b200.frames for larger frames)')
    return np.ascontiguousarray(a, dtype=np.uint8)


def load_png(path: str) -> np.ndarray:
    """PIL decode exactly like predict.py:55 -> uint8 [512,512,3]."""
    from PIL import Image
    with Image.open(path) as im:  # predict.py:55
        a = np.asarray(im)
    return _check_image(a, path)


def decode_png(path: str, out: Optional[np.ndarray] = None) -> np.ndarray:
    """Thread-pool friendly decode (OpenCV when importable, else PIL): the same pixels as load_png (PNG is lossless;
    tests/test_host_cpu.py checks it on the reference's 12 inputs). Writes into `out` (e.g. a row of a pinned batch
    buffer) when given."""
    if _cv2 is not None:
        a = _cv2.imread(path, _cv2.IMREAD_UNCHANGED)
        if a is None:
            raise OSError(f'cannot read {path}')
        if a.dtype != np.uint8:
            a = (a >> 8).astype(np.uint8) if a.dtype == np.uint16 else a.astype(np.uint8)
        if a.ndim == 3 and a.shape[2] >= 3:
            a = _cv2.cvtColor(a, _cv2.COLOR_BGRA2RGB if a.shape[2] == 4 else _cv2.COLOR_BGR2RGB)
        a = _check_image(a, path)
    else:
        a = load_png(path)
    if out is not None:
        np.copyto(out, a)
        return out
    return a


def encode_png(img_u8: np.ndarray, path: str, compress_level: Optional[int] = None):
    """uint8 [h,w,3] -> PNG file. compress_level None = the encoder's default (what predict.py:70's Image.save uses
    with PIL); the pixels are identical at every level."""
    if _cv2 is not None:
        params = [] if compress_level is None else [_cv2.IMWRITE_PNG_COMPRESSION, int(compress_level)]
        ok, buf = _cv2.imencode('.png', _cv2.cvtColor(img_u8, _cv2.COLOR_RGB2BGR), params)
        if not ok:
            raise OSError(f'cannot encode {path}')
        with open(path, 'wb') as f:
            f.write(buf.tobytes())
    else:
        from PIL import Image
        kw = {} if compress_level is None else {'compress_level': int(compress_level)}
        Image.fromarray(img_u8).save(path, **kw)


def run(model: Model, files: List[str], output_dir: str, batch: int = 1, faithful: bool = False,
        prefix: str = '', verbose: bool = True) -> List[str]:
    """The reference loop as written (predict.py:50-72): per chunk, blocking decode -> predict -> blocking encode."""
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


def run_pipelined(model: Model, files: List[str], output_dir: str, batch: int = 16, workers: Optional[int] = None,
                  prefix: str = '', compress_level: Optional[int] = 1, stats: Optional[Dict[str, float]] = None) -> List[str]:
    """predict.py:50-72 as a three-stage pipeline: a thread pool decodes chunk k+1 straight into a pinned
    batch buffer and encodes chunk k-1 out of another one while the GPU runs chunk k through itmo_predict_host (whose
    own H2D / compute / D2H stages are double-buffered inside the library). Two pinned input and two pinned output
    buffers; the same pixels as run(). `stats` (optional dict) receives files/s and where the wall time went."""
    import time
    from concurrent.futures import ThreadPoolExecutor, wait
    from .model import pinned_empty
    os.makedirs(output_dir, exist_ok=True)
    workers = workers or min(32, os.cpu_count() or 4)
    batch = max(1, min(batch, max(1, len(files))))
    chunks = [files[i:i + batch] for i in range(0, len(files), batch)]
    shape = (batch, graph.IMG, graph.IMG, 3)
    bin_ = [pinned_empty(shape, np.uint8) for _ in range(2)]
    bout = [pinned_empty(shape, np.uint8) for _ in range(2)]
    written: List[str] = []
    t_wait_dec = t_gpu = t_wait_enc = 0.0
    t_start = time.perf_counter()
    with ThreadPoolExecutor(workers) as pool:
        def submit_decode(k):
            return [pool.submit(decode_png, p, bin_[k % 2][i]) for i, p in enumerate(chunks[k])]
        dec = submit_decode(0) if chunks else []
        enc_pending = [[], []]
        for k, chunk in enumerate(chunks):
            t0 = time.perf_counter()
            for f in dec:
                f.result()                                  # chunk k decoded (raises the worker's error, if any)
            t1 = time.perf_counter()
            if k + 1 < len(chunks):
                dec = submit_decode(k + 1)                  # decode of k+1 overlaps the GPU call below
            wait(enc_pending[k % 2])                        # bout[k % 2] fully encoded (chunk k-2)
            for f in enc_pending[k % 2]:
                f.result()
            t2 = time.perf_counter()
            n = len(chunk)
            model.predict_u8(bin_[k % 2][:n], out=bout[k % 2][:n])   # releases the GIL: the pool keeps working
            t3 = time.perf_counter()
            futs = []
            for i, p in enumerate(chunk):
                dst = os.path.join(output_dir, prefix + os.path.basename(p))
                futs.append(pool.submit(encode_png, bout[k % 2][i], dst, compress_level))
                written.append(dst)
            enc_pending[k % 2] = futs
            t_wait_dec += t1 - t0
            t_wait_enc += t2 - t1
            t_gpu += t3 - t2
        t0 = time.perf_counter()
        for futs in enc_pending:
            for f in futs:
                f.result()
        t_wait_enc += time.perf_counter() - t0
    total = time.perf_counter() - t_start
    if stats is not None:
        stats.update(files=len(files), seconds=total, files_per_s=len(files) / total if total > 0 else 0.0,
                     gpu_seconds=t_gpu, waiting_for_decode_seconds=t_wait_dec, waiting_for_encode_seconds=t_wait_enc,
                     codec_share_of_wall=(t_wait_dec + t_wait_enc) / total if total > 0 else 0.0, workers=workers,
                     batch=batch, codec='opencv' if _cv2 is not None else 'pil')
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
    ap.add_argument('--workers', type=int, default=None, help='codec threads of the pipelined path (default: all cores)')
    ap.add_argument('--no-pipeline', action='store_true', help='blocking decode -> predict -> encode like the reference')
    a = ap.parse_args(argv)
    path = a.model or checkpoint.find_reference_checkpoint()
    if path is None:
        print(f'error: {checkpoint.REFERENCE_CHECKPOINT} not found (set --model or $ITMO_CKPT)', file=sys.stderr)
        return 2
    model = load_model(path, precision=a.precision, device=a.device, max_batch=max(1, a.batch))
    files = list_inputs(a.input_dir)
    if a.faithful or a.no_pipeline or a.batch <= 1:
        run(model, files, a.output_dir, batch=max(1, a.batch), faithful=a.faithful)
    else:
        st: Dict[str, float] = {}
        run_pipelined(model, files, a.output_dir, batch=a.batch, workers=a.workers, stats=st)
        print(f"{st['files']} files in {st['seconds']:.2f} s = {st['files_per_s']:.1f} files/s "
              f"(GPU calls {st['gpu_seconds']:.2f} s, waiting for the {st['codec']} codec "
              f"{100 * st['codec_share_of_wall']:.0f} % of the wall time, {st['workers']} threads)")
    return 0


if __name__ == '__main__':
    sys.exit(main())
