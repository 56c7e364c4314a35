"""One training micro-batch on the GPU with per-layer comparison against the oracle (debug aid).
usage: python tools/train_debug.py [n] [h] [w] [--no-ref]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import cnn_itmo_b200 as M
from cnn_itmo_b200 import train
from oracle import unet_ref as R, unet_train_ref as T

args = [a for a in sys.argv[1:] if not a.startswith('--')]
n, h, w = (int(args[0]) if args else 2), (int(args[1]) if len(args) > 1 else 256), (int(args[2]) if len(args) > 2 else 128)
ws = M.initial_weights(3) if '--init' in sys.argv else R.synthetic_weights(3)
x, y = T.synthetic_pair(n, h, w, seed=11)
tr = train.Trainer(ws, micro_batch=n, image_hw=(h, w))
dev = torch.device('cuda:0')
tr.zero_grad()
tr.backward(torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev), float(y.size), seed=0, dropout=False, freeze_moving=True)
torch.cuda.synchronize()
sq, hits = tr.metrics()
print('gpu sq_err_sum', sq, 'hits', hits)
if '--no-ref' in sys.argv:
    sys.exit(0)
ref = T.train_forward_backward(ws, x.astype(np.float32) / 255.0, y, None, return_acts=True, emulate_bf16='--emu' in sys.argv)
print('ref sq_err_sum', ref['sq_err_sum'], 'acc', ref['acc'])
rel = lambda a, b: float(np.linalg.norm(np.asarray(a, np.float64).ravel() - np.asarray(b, np.float64).ravel()) / max(np.linalg.norm(np.asarray(b, np.float64).ravel()), 1e-30))
for li in range(18):
    print(f'layer {li:2d} z {rel(tr.read_activation(li, "z", n), ref["z"][li]):.3e}  a {rel(tr.read_activation(li, "a", n), ref["a"][li]):.3e}')
g = tr.get_grads()
for li in range(19):
    print(f'layer {li:2d} ' + '  '.join(f'{k} {rel(g[li][k], ref["grads"][li][k]):.3e}' for k in ('kernel', 'bias', 'gamma', 'beta') if k in g[li]))
