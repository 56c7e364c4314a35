import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from cnn_itmo_b200 import _lib as L
lib = L.load()
dev = torch.device('cuda:0')
shapes = [(32, 32, 1, 64, 32), (64, 64, 1, 128, 64), (32, 64, 1, 256, 128), (128, 128, 2, 64, 64), (256, 128, 1, 64, 32)]
if len(sys.argv) > 1:
    shapes = [tuple(int(v) for v in sys.argv[1:6])]
for cin, cout, n, h, w in shapes:
    g = torch.Generator().manual_seed(1)
    x = torch.randn(n, h, w, cin, generator=g).to(torch.bfloat16)
    gy = torch.randn(n, h, w, cout, generator=g).to(torch.bfloat16)
    xd, gd = x.to(dev), gy.to(dev)
    dw = torch.zeros(3, 3, cin, cout, dtype=torch.float32, device=dev)
    try:
        L.check(lib.itmo_wgrad3x3(xd.data_ptr(), cin, gd.data_ptr(), cout, n, h, w, dw.data_ptr(), None))
        torch.cuda.synchronize()
    except Exception as e:  # noqa
        print((cin, cout, n, h, w), 'FAILED', str(e)[:200]); break
    k = torch.zeros(cout, cin, 3, 3, dtype=torch.float32, requires_grad=True)
    y = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), k, padding=1)
    y.backward(gy.float().permute(0, 3, 1, 2))
    ref = k.grad.permute(2, 3, 1, 0).numpy().astype(np.float64)
    got = dw.cpu().numpy().astype(np.float64)
    print((cin, cout, n, h, w), 'rel', np.linalg.norm(got - ref) / np.linalg.norm(ref), flush=True)
