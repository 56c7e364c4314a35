"""Regenerates the committed fixtures under tests/golden/ (run in the build container, where the reference
tree is mounted read-only at /root/reference; the GPU box never sees that path).

  layers_table.json     parsed rows of the reference's layers.txt (name, class, shape, params, inbound) + totals:
                        the only artefact of the reference that pins the U-Net's topology / parameter counts
  images_to_predict/    the reference's 12 input PNGs and the 12 outputs its authors committed
                        (images_to_predict/{input,output}); the outputs are golden vectors ONLY together with
                        the stripped checkpoint saved7-model-218-0.73.hdf5 (+-1 LSB, predict.py:64 truncates)
  oracle_small.npz      seeded synthetic weights -> oracle (numpy float64) outputs on a 32x32 input, so the
                        oracle itself cannot drift silently
"""
import json
import os
import re
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = '/root/reference'


def parse_layers_txt(path):
    rows, totals = [], {}
    cur = None
    for line in open(path):
        line = line.rstrip('\n')
        if not line or set(line) <= {'_', '='}:
            continue
        m = re.match(r'^(Total|Trainable|Non-trainable) params: ([\d,]+)', line)
        if m:
            totals[m.group(1).lower().replace('-', '_')] = int(m.group(2).replace(',', ''))
            continue
        if line.startswith('Layer (type)'):
            continue
        name_cls, shape, params, conn = line[:32].strip(), line[32:53].strip(), line[53:65].strip(), line[65:].strip()
        if name_cls:
            n, c = name_cls.split(' ', 1)
            cur = {'name': n, 'cls_prefix': c.strip('()'), 'shape': shape, 'params': int(params), 'inbound': []}
            rows.append(cur)
        if conn:
            cur['inbound'].append(conn)
    return {'rows': rows, 'totals': totals}


def main():
    tab = parse_layers_txt(os.path.join(REF, 'layers.txt'))
    json.dump(tab, open(os.path.join(HERE, 'layers_table.json'), 'w'), indent=1)
    for sub in ('input', 'output'):
        dst = os.path.join(HERE, 'images_to_predict', sub)
        os.makedirs(dst, exist_ok=True)
        for f in sorted(os.listdir(os.path.join(REF, 'images_to_predict', sub))):
            shutil.copyfile(os.path.join(REF, 'images_to_predict', sub, f), os.path.join(dst, f))
    from oracle import unet_ref as R
    ws = R.synthetic_weights(218)
    x = np.random.default_rng(7).random((1, 32, 32, 3))
    outs, pooled = R.forward_np(ws, x, return_all=True)
    np.savez_compressed(os.path.join(HERE, 'oracle_small.npz'), x=x, y=outs[18], l1=outs[1], p1=pooled[1], l10=outs[10],
                        l17=outs[17])
    print('golden fixtures written:', len(tab['rows']), 'layer rows', tab['totals'])


if __name__ == '__main__':
    main()
