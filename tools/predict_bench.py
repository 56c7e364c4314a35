#!/usr/bin/env python
"""Host-pipeline numbers of the predict.py flow on one B200; prints ONE JSON object.

  files/s of the reference's own loop (blocking decode -> predict -> encode, one image per call, predict.py:50-72) and of
  the pipelined path (thread-pool PNG codec around itmo_predict_host) on the 12 reference input PNGs x REPEAT, with the
  codec's share of the wall time, and the latency of one batch-1 predict call with and without CUDA-graph replay of the
  18-launch forward.   usage: python tools/predict_bench.py [--repeat 100] [--batch 64] [--out profiles/x.json]
"""
import argparse
import json
import os
import shutil
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--repeat', type=int, default=100)
    ap.add_argument('--batch', type=int, default=64)
    ap.add_argument('--blocking-files', type=int, default=120, help='files of the blocking reference loop (bounded)')
    ap.add_argument('--out', default=None)
    a = ap.parse_args()
    import numpy as np
    import __graft_entry__ as g
    g.build()
    import cnn_itmo_b200 as M
    from cnn_itmo_b200 import predict as P
    import bench

    class HZ:  # the little of bench.py's harness calibrated_checkpoint needs
        rank = 0
        local_rank = 0

        @staticmethod
        def barrier():
            pass
    ckpt = bench.calibrated_checkpoint(HZ)
    src = os.path.join(ROOT, 'tests', 'golden', 'images_to_predict', 'input')
    pngs = sorted(os.path.join(src, f) for f in os.listdir(src) if f.lower().endswith('.png'))
    tmp = tempfile.mkdtemp(prefix='itmo_predict_bench_')
    ind, outd = os.path.join(tmp, 'in'), os.path.join(tmp, 'out')
    os.makedirs(ind)
    files = []
    for r in range(a.repeat):
        for p in pngs:
            dst = os.path.join(ind, f'{r:03d}_{os.path.basename(p)}')
            os.symlink(p, dst)
            files.append(dst)
    res = {'files': len(files), 'inputs': f'{len(pngs)} reference PNGs x {a.repeat}', 'host_cores': os.cpu_count()}

    model = M.load_model(ckpt, max_batch=a.batch)
    res['precision'] = model.precision
    # warm-up (kernel attributes, graphs, page faults of the pinned buffers)
    P.run_pipelined(model, files[:2 * a.batch], outd, batch=a.batch)
    st = {}
    P.run_pipelined(model, files, outd, batch=a.batch, stats=st)
    res['pipelined'] = {k: (round(v, 4) if isinstance(v, float) else v) for k, v in st.items()}
    # the reference's loop as written: one image per predict call, blocking PIL codec
    m1 = M.load_model(ckpt, max_batch=1)
    nb = min(a.blocking_files, len(files))
    P.run(m1, files[:4], outd, batch=1, verbose=False)
    t0 = time.perf_counter()
    P.run(m1, files[:nb], outd, batch=1, verbose=False)
    dt = time.perf_counter() - t0
    res['blocking_reference_loop'] = {'files': nb, 'seconds': round(dt, 3), 'files_per_s': round(nb / dt, 2)}
    # batch-1 latency, eager 18 launches vs one graph replay
    img = np.ascontiguousarray(P.load_png(pngs[0])[None])
    lat = {}
    for name, flag in (('eager', False), ('graph', True)):
        mm = M.load_model(ckpt, max_batch=1)
        mm.use_cuda_graph = flag          # read when the device context is created (first call)
        for _ in range(20):
            mm.predict_u8(img)
        ts = []
        for _ in range(200):
            t0 = time.perf_counter()
            mm.predict_u8(img)
            ts.append(time.perf_counter() - t0)
        ts.sort()
        lat[name] = {'median_ms': round(1e3 * ts[len(ts) // 2], 4), 'p10_ms': round(1e3 * ts[len(ts) // 10], 4),
                     'p90_ms': round(1e3 * ts[9 * len(ts) // 10], 4)}
        mm.close()
    res['batch1_predict_u8_latency'] = lat
    shutil.rmtree(tmp, ignore_errors=True)
    line = json.dumps(res)
    print(line)
    if a.out:
        with open(a.out, 'w') as f:
            f.write(line + '\n')


if __name__ == '__main__':
    main()
