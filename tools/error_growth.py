"""Per-layer error growth of the reduced-precision modes on the calibrated checkpoint (run under gpurun).

For every stored activation of the forward pass (itmo_read_activation) the relative L2 error and the PSNR (peak = the
oracle tensor's max-abs) against the fp32 CPU oracle, for bf16 / fp16 / tf32 / fp32 storage — so that the loss of the
16-bit modes is LOCATED layer by layer, not guessed. Writes gpurun_out/error_growth.{json,md}; the summary is kept
under profiles/.

    python tools/error_growth.py [--images 2]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--images', type=int, default=2)
    ap.add_argument('--out', default=os.path.join(ROOT, 'gpurun_out', 'error_growth'))
    a = ap.parse_args()
    import cnn_itmo_b200 as M
    from oracle import unet_ref as R
    ws = R.calibrated_weights(218)
    imgs = R.synthetic_images_u8(a.images, seed=5)
    x = R.preprocess_u8(imgs)
    outs, pooled = R.forward_torch(R.TorchWeights(ws), x, return_all=True)
    rows = {}
    precs = ['bf16', 'fp16', 'tf32', 'fp32']
    for prec in precs:
        m = M.Model(ws, precision=prec, max_batch=a.images)
        final = m.predict(x)
        per = []
        for li in range(17):
            act = m.read_activation(li).astype(np.float64)
            ref = outs[li].astype(np.float64)
            rel = float(np.linalg.norm(act - ref) / np.linalg.norm(ref))
            per.append({'layer': R.LAYERS[li].keras_name, 'rel_l2': rel, 'psnr_db': R.psnr(act, ref, peak=float(np.abs(ref).max())),
                        'ref_absmax': float(np.abs(ref).max()), 'dev_absmax': float(np.abs(act).max())})
        ref = outs[18]
        per.append({'layer': 'conv2d_15 (sigmoid out)', 'rel_l2': float(np.linalg.norm(final - ref) / np.linalg.norm(ref)),
                    'psnr_db': R.psnr(final, ref), 'ref_absmax': float(ref.max()), 'dev_absmax': float(final.max()),
                    'max_abs': float(np.abs(final - ref).max())})
        rows[prec] = per
        m.close()
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out + '.json', 'w') as f:
        json.dump({'images': a.images, 'weights': 'calibrated(218)', 'rows': rows}, f, indent=1)
    lines = ['| layer | oracle max-abs | ' + ' | '.join(f'{p} rel-L2 | {p} PSNR dB' for p in precs) + ' |',
             '|---|---|' + '---|---|' * len(precs)]
    for i in range(18):
        cells = []
        for p in precs:
            r = rows[p][i]
            cells += [f"{r['rel_l2']:.2e}", f"{r['psnr_db']:.1f}"]
        lines.append(f"| {rows[precs[0]][i]['layer']} | {rows[precs[0]][i]['ref_absmax']:.3g} | " + ' | '.join(cells) + ' |')
    md = '\n'.join(lines) + '\n'
    with open(a.out + '.md', 'w') as f:
        f.write(md)
    print(md)


if __name__ == '__main__':
    main()
