"""GPU bring-up diagnostics for the per-layer operators (run under gpurun; not part of the test suite).

python tools/gpu_diag.py            -> runs every case in its own subprocess (a trapped kernel kills only
                                       its case), prints one JSON line per case, writes gpurun_out/diag.jsonl
python tools/gpu_diag.py --case i   -> one case in-process
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# (name, prec, kind, n, h, w, c0, c1, cout, pool, variant)
CASES = []
for prec in ('bf16', 'tf32'):
    CASES += [
        (f'{prec}_tiny_center', prec, 'conv', 1, 16, 16, 64, 0, 64, False, 'center'),
        (f'{prec}_tiny', prec, 'conv', 1, 16, 16, 64, 0, 64, False, 'rand'),
        (f'{prec}_small_k2', prec, 'conv', 2, 16, 32, 128, 0, 128, False, 'rand'),
        (f'{prec}_small_pool', prec, 'conv', 2, 16, 32, 64, 0, 64, True, 'rand'),
        (f'{prec}_small_2src', prec, 'conv', 1, 16, 32, 64, 128, 128, False, 'rand'),
        (f'{prec}_small_n256', prec, 'conv', 1, 16, 16, 128, 0, 256, False, 'rand'),
        (f'{prec}_small_n512', prec, 'conv', 1, 16, 16, 256, 0, 512, False, 'rand'),
        (f'{prec}_small_c32', prec, 'conv', 1, 16, 32, 32, 0, 32, True, 'rand'),
        (f'{prec}_small_c32_64', prec, 'conv', 1, 16, 32, 32, 64, 64, False, 'rand'),
        (f'{prec}_convT_small', prec, 'convT', 1, 16, 16, 128, 0, 64, False, 'rand'),
        (f'{prec}_convT_512', prec, 'convT', 1, 16, 16, 512, 0, 512, False, 'rand'),
        (f'{prec}_head', prec, 'head', 1, 16, 32, 32, 64, 64, False, 'rand'),
        (f'{prec}_rgb_u8', prec, 'rgb', 2, 32, 32, 3, 0, 32, False, 'u8'),
        (f'{prec}_rgb_f32', prec, 'rgb', 2, 32, 32, 3, 0, 32, False, 'f32'),
    ]
# full-size layers, timing
for prec in ('bf16', 'tf32'):
    CASES += [
        (f'{prec}_L11', prec, 'conv', 8, 64, 64, 256, 512, 512, False, 'rand'),
        (f'{prec}_L13', prec, 'conv', 8, 128, 128, 128, 256, 256, False, 'rand'),
        (f'{prec}_L15', prec, 'conv', 8, 256, 256, 64, 128, 128, False, 'rand'),
        (f'{prec}_L17', prec, 'conv', 8, 512, 512, 32, 64, 64, False, 'rand'),
        (f'{prec}_L1', prec, 'conv', 8, 512, 512, 32, 0, 32, True, 'rand'),
        (f'{prec}_L16T', prec, 'convT', 8, 256, 256, 128, 0, 64, False, 'rand'),
        (f'{prec}_L10T', prec, 'convT', 8, 32, 32, 512, 0, 512, False, 'rand'),
    ]


def run_case(idx):
    import numpy as np
    import torch
    import importlib
    lib_mod = importlib.import_module('cnn_itmo_b200._lib')
    lib = lib_mod.load()
    name, prec, kind, n, h, w, c0, c1, cout, pool, variant = CASES[idx]
    P = lib_mod.PREC_BF16 if prec == 'bf16' else lib_mod.PREC_TF32
    dt = torch.bfloat16 if prec == 'bf16' else torch.float32
    g = torch.Generator().manual_seed(1000 + idx)
    dev = 'cuda:0'

    def quant(t):
        if prec == 'bf16':
            return t.to(torch.bfloat16).float()
        return (t.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)

    res = {'case': name}
    F = torch.nn.functional
    bias = torch.rand(cout, generator=g) * 0.2 - 0.1
    scale = torch.randn(cout, generator=g) * 0.25 + 1.0
    scale[::7] *= -1
    shift = torch.randn(cout, generator=g) * 0.1
    if kind == 'rgb':
        if variant == 'u8':
            xin = torch.randint(0, 256, (n, h, w, 3), generator=g, dtype=torch.uint8)
            xf = xin.float() / 255.0
        else:
            xin = torch.rand(n, h, w, 3, generator=g)
            xf = xin
        k = torch.randn(3, 3, 3, cout, generator=g) * (2.0 / 27) ** 0.5
        ref = F.conv2d(xf.permute(0, 3, 1, 2).double(), k.permute(3, 2, 0, 1).double(), bias.double(), padding=1)
        ref = torch.relu(ref) * scale.double()[None, :, None, None] + shift.double()[None, :, None, None]
        ref = ref.permute(0, 2, 3, 1)
        out = torch.empty(n, h, w, cout, dtype=dt, device=dev)
        d = lambda t: t.to(dev).contiguous()
        xin_d, k_d, b_d, s_d, t_d = d(xin), d(k), d(bias), d(scale), d(shift)
        lib_mod.check(lib.itmo_conv3x3_bn_rgb(P, xin_d.data_ptr(), lib_mod.DT_U8 if variant == 'u8' else lib_mod.DT_F32,
                                              n, h, w, k_d.data_ptr(), b_d.data_ptr(), s_d.data_ptr(), t_d.data_ptr(),
                                              cout, out.data_ptr(), None))
        torch.cuda.synchronize()
        err = (out.float().cpu().double() - ref).abs()
        res.update(max_abs=float(err.max()), ref_max=float(ref.abs().max()))
        return res

    cin = c0 + c1
    x0 = quant(torch.randn(n, h, w, c0, generator=g))
    x1 = quant(torch.randn(n, h, w, c1, generator=g)) if c1 else None
    xcat = torch.cat([x0, x1], dim=3) if c1 else x0
    if kind == 'convT':
        k = quant(torch.randn(2, 2, cout, cin, generator=g) * (2.0 / (4 * cin)) ** 0.5)
        ref = F.conv_transpose2d(xcat.permute(0, 3, 1, 2).double(), k.permute(3, 2, 0, 1).double(), bias.double(), stride=2)
    else:
        k = quant(torch.randn(3, 3, cin, cout, generator=g) * (2.0 / (9 * cin)) ** 0.5)
        if variant == 'center':
            m = torch.zeros(3, 3, 1, 1)
            m[1, 1] = 1
            k = k * m
        ref = F.conv2d(xcat.permute(0, 3, 1, 2).double(), k.permute(3, 2, 0, 1).double(), bias.double(), padding=1)
    ref = torch.relu(ref) * scale.double()[None, :, None, None] + shift.double()[None, :, None, None]
    ref_pool = F.max_pool2d(ref, 2, 2).permute(0, 2, 3, 1) if pool else None
    ref = ref.permute(0, 2, 3, 1).contiguous()
    d = lambda t: t.to(dev).contiguous()
    x0_d = d(x0.to(dt))
    x1_d = d(x1.to(dt)) if c1 else None
    k_d, b_d, s_d, t_d = d(k), d(bias), d(scale), d(shift)
    ho, wo = (2 * h, 2 * w) if kind == 'convT' else (h, w)
    out = torch.full((n, ho, wo, cout), float('nan'), dtype=dt, device=dev)
    pooled = torch.full((n, h // 2, w // 2, cout), float('nan'), dtype=dt, device=dev) if pool else None

    def launch():
        if kind == 'conv':
            return lib.itmo_conv3x3_bn(P, x0_d.data_ptr(), c0, x1_d.data_ptr() if c1 else None, c1, n, h, w,
                                       k_d.data_ptr(), b_d.data_ptr(), s_d.data_ptr(), t_d.data_ptr(), cout,
                                       out.data_ptr(), pooled.data_ptr() if pool else None, None)
        if kind == 'convT':
            return lib.itmo_convT2x2_bn(P, x0_d.data_ptr(), cin, n, h, w, k_d.data_ptr(), b_d.data_ptr(),
                                        s_d.data_ptr(), t_d.data_ptr(), cout, out.data_ptr(), None)
        raise AssertionError

    if kind == 'head':
        hk = torch.randn(1, 1, 64, 3, generator=g) * 0.2
        hb = torch.randn(3, generator=g) * 0.1
        refh = torch.sigmoid(torch.einsum('nhwc,co->nhwo', ref, hk[0, 0].double()) + hb.double())
        hk_d, hb_d = d(hk), d(hb)
        for odt, tdt in ((lib_mod.DT_F32, torch.float32), (lib_mod.DT_U8, torch.uint8)):
            o = torch.zeros(n, h, w, 3, dtype=tdt, device=dev)
            lib_mod.check(lib.itmo_conv3x3_bn_head(P, x0_d.data_ptr(), c0, x1_d.data_ptr() if c1 else None, c1, n, h, w,
                                                   k_d.data_ptr(), b_d.data_ptr(), s_d.data_ptr(), t_d.data_ptr(),
                                                   hk_d.data_ptr(), hb_d.data_ptr(), o.data_ptr(), odt, None))
            torch.cuda.synchronize()
            if tdt == torch.float32:
                res['head_f32_max_abs'] = float((o.cpu().double() - refh).abs().max())
            else:
                want = (refh.float() * 255).to(torch.uint8)
                res['head_u8_max_abs'] = int((o.cpu().int() - want.int()).abs().max())
        return res

    lib_mod.check(launch())
    torch.cuda.synchronize()
    o = out.float().cpu().double()
    err = (o - ref).abs()
    res.update(max_abs=float(err.max()), mean_abs=float(err.mean()), ref_max=float(ref.abs().max()),
               nan=int(torch.isnan(o).sum()))
    if float(err.max()) > 0.05 or res['nan']:
        # localise: error per output row / col / channel block
        e = torch.nan_to_num(err, nan=1e3)
        res['err_by_row'] = [round(float(v), 3) for v in e.amax(dim=(0, 2, 3))[:16]]
        res['err_by_col'] = [round(float(v), 3) for v in e.amax(dim=(0, 1, 3))[:32]]
        res['err_by_ch16'] = [round(float(v), 3) for v in e.amax(dim=(0, 1, 2)).view(-1, 16).amax(dim=1)[:32]]
    if pool:
        pe = (pooled.float().cpu().double() - ref_pool).abs()
        res['pool_max_abs'] = float(torch.nan_to_num(pe, nan=1e3).max())
    if n >= 8:  # timing of the full-size layers
        for _ in range(3):
            lib_mod.check(launch())
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 10
        ev0.record()
        for _ in range(reps):
            lib_mod.check(launch())
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1) / reps
        taps = 4 if kind == 'convT' else 9
        flops = 2.0 * n * h * w * taps * cin * cout
        res.update(ms=round(ms, 4), tflops=round(flops / ms / 1e9, 1))
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--case', type=int, default=-1)
    ap.add_argument('--only', type=str, default='')
    a = ap.parse_args()
    if a.case >= 0:
        print('DIAG ' + json.dumps(run_case(a.case)), flush=True)
        return
    os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
    with open(os.path.join(ROOT, 'gpurun_out', 'diag.jsonl'), 'w') as f:
        for i, c in enumerate(CASES):
            if a.only and a.only not in c[0]:
                continue
            t0 = time.time()
            try:
                p = subprocess.run([sys.executable, os.path.abspath(__file__), '--case', str(i)], capture_output=True,
                                   text=True, timeout=180)
                line = [l for l in p.stdout.splitlines() if l.startswith('DIAG ')]
                if line:
                    rec = json.loads(line[-1][5:])
                else:
                    rec = {'case': c[0], 'error': (p.stdout[-800:] + p.stderr[-1500:])}
            except subprocess.TimeoutExpired:
                rec = {'case': c[0], 'error': 'timeout'}
            rec['wall_s'] = round(time.time() - t0, 1)
            print(json.dumps(rec), flush=True)
            f.write(json.dumps(rec) + '\n')
            f.flush()


if __name__ == '__main__':
    main()
