#!/usr/bin/env python
"""bench.py — throughput of the CNN-ITMO U-Net hot path (predict.py:62, model.py:204-278) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's sm_100a path
    python bench.py --impl reference [--steps K] [--warmup W]      # CPU restatement of the reference path
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N   # one rank per GPU (weak scaling)

A "step" is one forward pass of the whole U-Net over one batch of synthetic 512x512 SDR images per GPU
(BASELINE.json configs[2]: full U-Net inference, batch 256, bf16 operands / fp32 accumulate). Weights are
random-init (he_normal, exactly what ``U_net()`` creates): the committed checkpoint the reference names is
absent from the reference tree (.MISSING_LARGE_BLOBS). Rank 0 prints ONE JSON line:

  value      images/s over all ranks, inputs already resident in HBM (itmo_forward on device buffers)
  e2e        images/s through Model.predict_u8 -> itmo_predict_host: pinned HOST uint8 in, HOST uint8 out,
             H2D and D2H copies inside the timed region
  roofline   the implicit-GEMM conv kernel (17 launches per arena chunk = 99.7 % of the FLOPs): algorithmic
             FLOPs / summed CUDA-event duration of those launches inside the timed region, against the
             measured bf16 tensor peak of MEASURED_PEAKS.json
  cpu_baseline  the torch-CPU (oneDNN) restatement of the reference graph (oracle/unet_ref.py) timed on this
             box's host cores on a bounded sample — TensorFlow itself cannot be installed here

L2 hygiene: every step streams > 40 GB of activations through a 126 MB L2 (inputs alone are 201 MB), so no
explicit flush is needed; ``config.l2`` says so.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

IMG = 512
MPIX_PER_IMAGE = IMG * IMG / 1e6
METRIC = 'U-Net SDR->HDR inference throughput (512x512 images)'
UNIT = 'images/s'


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peaks():
    """Roofline denominators: MEASURED_PEAKS.json (driver-written) or the profiling guide's fallback."""
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    try:
        with open(p) as f:
            d = json.load(f)
        return {'hbm_gbs': float(d['hbm_gbs']), 'tflops_burst': float(d['bf16_tflops']),
                'tflops_sustained': float(d.get('bf16_tflops_sustained', d['bf16_tflops'])), 'source': 'measured'}
    except Exception:  # noqa: BLE001
        return {'hbm_gbs': 6650.0, 'tflops_burst': 1590.0, 'tflops_sustained': 1400.0, 'source': 'fallback'}


# --------------------------------------------------------------------------------------------------
# clocks: sample nvidia-smi DURING the timed region
# --------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
              'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
              'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(prefix='itmo_clocks_', suffix='.csv')
            os.close(fd)
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.gpu), f'--query-gpu={self.FIELDS}',
                                          '--format=csv,noheader,nounits', '-lms', '50'],
                                         stdout=open(self.path, 'w'), stderr=subprocess.DEVNULL)
        except Exception:  # noqa: BLE001
            self.proc = None

    def stop(self):
        out = {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': [], 'samples': 0}
        if self.proc is None:
            return out
        try:
            self.proc.terminate()
            self.proc.wait(timeout=5)
        except Exception:  # noqa: BLE001
            try:
                self.proc.kill()
            except Exception:  # noqa: BLE001
                pass
        sm, mx, pw, reasons = [], [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        try:
            with open(self.path) as f:
                for line in f:
                    parts = [p.strip() for p in line.split(',')]
                    if len(parts) < 9:
                        continue
                    try:
                        sm.append(float(parts[1]))
                        mx.append(float(parts[2]))
                        pw.append(float(parts[3]))
                    except ValueError:
                        continue
                    for nm, v in zip(names, parts[5:9]):
                        if v.lower().startswith('active'):
                            reasons.add(nm)
            os.unlink(self.path)
        except Exception:  # noqa: BLE001
            pass
        if sm:
            # "under load" = samples whose power is in the upper half of what was seen
            thr = 0.5 * max(pw) if pw else 0
            load = sorted(s for s, p in zip(sm, pw) if p >= thr) or sorted(sm)
            out.update(sm_mhz=load[len(load) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm),
                       power_w_max=max(pw) if pw else None)
        return out


# --------------------------------------------------------------------------------------------------
# CPU restatement (oracle) legs — the only places bench.py touches oracle/
# --------------------------------------------------------------------------------------------------
def cpu_reference_time(weights, n_images_hint: int, budget_s: float = 15.0):
    """Times the torch-CPU restatement of predict.py:50-64 (one image per call, like the reference loop) on all
    host cores. Returns (images_per_s, n_images, cores, seconds)."""
    import numpy as np
    import torch
    from oracle import unet_ref as R
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    tw = R.TorchWeights(weights)
    rng = np.random.default_rng(0)
    imgs = rng.integers(0, 256, (max(1, n_images_hint), IMG, IMG, 3), dtype=np.uint8)
    R.predict_like_reference(tw, imgs[:1])                      # warm-up (oneDNN primitive creation)
    t0 = time.perf_counter()
    R.predict_like_reference(tw, imgs[:1])
    t1 = time.perf_counter() - t0
    n = int(max(2, min(n_images_hint, budget_s / max(t1, 1e-3))))
    t0 = time.perf_counter()
    R.predict_like_reference(tw, imgs[:n])
    dt = time.perf_counter() - t0
    return n / dt, n, cores, dt


def run_reference_arm(a):
    """--impl reference: the reference's own CPU path cannot run here (TensorFlow 1.x / Keras 2.2 are not
    installable), so this times its CPU restatement (oracle/unet_ref.py, torch-CPU/oneDNN fp32) with all host
    threads. Each step = a bounded sample of the batch-256 workload, one image per predict call."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return 0
    import numpy as np
    import torch
    from oracle import unet_ref as R
    import cnn_itmo_b200 as M
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ws = M.initial_weights(seed=218)
    tw = R.TorchWeights(ws)
    a.cpu_sample = max(1, min(a.cpu_sample, 12))   # 20 default steps x 12 images ~ 1 minute of CPU work
    imgs = np.random.default_rng(0).integers(0, 256, (a.cpu_sample, IMG, IMG, 3), dtype=np.uint8)
    for _ in range(max(1, a.warmup)):
        R.predict_like_reference(tw, imgs[:1])
    t0 = time.perf_counter()
    for _ in range(a.steps):
        R.predict_like_reference(tw, imgs)
    dt = time.perf_counter() - t0
    ips = a.steps * a.cpu_sample / dt
    sample = (f'{a.cpu_sample} of the {a.batch} images per step, one image per predict call (predict.py:50-62), '
              f'torch-CPU fp32 restatement, {cores} threads')
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': round(ips, 4), 'unit': UNIT, 'n_gpus': a.gpus,
        'steps': a.steps, 'warmup': a.warmup, 'ms_per_step': round(1e3 * dt / a.steps, 3), 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'mpix_per_s': round(ips * MPIX_PER_IMAGE, 4),
        'config': workload_config(a, chunk=None),
        'cpu_baseline': {'value': round(ips, 4), 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': round(ips, 4), 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
        'note': 'TensorFlow-CPU (the reference runtime) is unavailable on this box; this is its CPU restatement',
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(a, chunk):
    cfg = {'workload': f'full U-Net forward (19 weight layers, 165.37 GFLOP/image), batch {a.batch} per GPU, '
                       f'512x512x3 uint8 in -> uint8 out, random-init weights (BASELINE configs[2])',
           'batch_per_gpu': a.batch, 'global_batch': a.batch * a.gpus, 'precision': a.precision,
           'parallelism': f'dp{a.gpus} (independent image batches, no collective on the data path)',
           'l2': 'inputs and activations far larger than L2 (>40 GB streamed per step); no explicit flush'}
    if chunk is not None:
        cfg['arena_chunk_images'] = chunk
    return cfg


# --------------------------------------------------------------------------------------------------
# the B200 arm
# --------------------------------------------------------------------------------------------------
def run_b200_arm(a):
    import numpy as np
    import torch
    import torch.distributed as dist

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device — the sm_100a library is the only implementation '
                         '(use --impl reference for the CPU restatement)')
    if world != a.gpus:
        if world == 1 and a.gpus > 1:
            raise SystemExit(f'bench.py --gpus {a.gpus}: launch with torchrun --nproc-per-node {a.gpus} '
                             '(one rank per GPU)')
        a.gpus = world
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)

    import __graft_entry__ as G
    if not os.path.exists(G.LIB):
        G.build()
    import cnn_itmo_b200 as M
    from cnn_itmo_b200.model import pinned_empty

    peaks = measured_peaks()
    ws = M.initial_weights(seed=218)
    model = M.Model(ws, precision=a.precision, device=local_rank, max_batch=a.chunk)
    B = a.batch
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    x = torch.randint(0, 256, (B, IMG, IMG, 3), dtype=torch.uint8, device=dev, generator=g)
    y = torch.empty_like(x)
    stream = torch.cuda.current_stream()

    def step():
        model.forward_device(x.data_ptr(), 'u8', y.data_ptr(), 'u8', B, stream=stream.cuda_stream)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    for _ in range(max(a.warmup, 3) if a.warmup else 0):
        step()
    barrier()
    model.set_profiling(True)
    model.layer_times(reset=True)
    launches0 = model.launch_count()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for _ in range(a.steps):
        step()
    ev1.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_total = ev0.elapsed_time(ev1)
    launches = model.launch_count() - launches0
    layer_times = model.layer_times(reset=True)
    model.set_profiling(False)
    if world > 1:
        t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / a.steps
    value = world * B * a.steps / (ms_total / 1e3)

    # sanity: the timed work really produced images (not NaN / constant)
    probe = y[0, ::64, ::64].float()
    if not bool(torch.isfinite(probe).all()):
        raise SystemExit('bench.py: non-finite output')

    # ---- end to end: pinned host uint8 -> predict_u8 -> pinned host uint8 --------------------------
    hx = pinned_empty((B, IMG, IMG, 3), np.uint8)
    hy = pinned_empty((B, IMG, IMG, 3), np.uint8)
    hx[...] = x.cpu().numpy()
    model.predict_u8(hx, out=hy)  # warm-up (allocates the device staging buffers)
    e2e_steps = max(1, min(a.steps, 5))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        model.predict_u8(hx, out=hy)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    e2e_value = world * B * e2e_steps / dt
    same = bool((torch.from_numpy(np.asarray(hy[:2])).to(dev) == y[:2]).all())
    if not same and not os.environ.get('ITMO_DBG'):  # ITMO_DBG timing experiments compute garbage on purpose
        raise SystemExit('bench.py: host path and device path disagree')

    # ---- roofline of the dominant kernel (conv_igemm_kernel), from the events of the timed region ---
    from cnn_itmo_b200 import graph
    per_layer = []
    igemm_flops = igemm_ms = 0.0
    for l in graph.LAYERS:
        ms, calls = layer_times[l.index]
        ks = graph.kernel_shape(l)
        if l.kind == 'convT':
            fl = 2.0 * (2 * l.hin) ** 2 * ks[2] * ks[3]
        else:
            fl = 2.0 * l.hin ** 2 * ks[0] * ks[1] * ks[2] * ks[3]
        if l.index == 17:
            fl += 2.0 * IMG * IMG * 64 * 3  # fused 1x1 head
        if l.kind == 'head':
            continue
        ms_per_step = ms / a.steps
        tf = fl * B / (ms_per_step * 1e-3) / 1e12 if ms_per_step > 0 else 0.0
        # algorithmic HBM bytes per image: every source read once, the output written once (+ the
        # pooled copy), the weights once per launch; uint8 pixels in front of conv2d_1, uint8 RGB behind conv2d_14
        es = 2 if a.precision in ('bf16', 'fp16') else 4
        hout = 2 * l.hin if l.kind == 'convT' else l.hin
        by = l.hin ** 2 * (3 if l.index == 0 else sum(l.cin) * es)
        by += IMG * IMG * 3 if l.index == 17 else hout ** 2 * l.cout * es * (1.25 if l.pool else 1.0)
        gbs = by * B / (ms_per_step * 1e-3) / 1e9 if ms_per_step > 0 else 0.0
        t_tensor = fl / (peaks['tflops_sustained'] * 1e12)
        t_hbm = by / (peaks['hbm_gbs'] * 1e9)
        per_layer.append({'layer': l.name, 'ms_per_step': round(ms_per_step, 4), 'tflops': round(tf, 1),
                          'gbs': round(gbs, 1), 'bound': 'tensor' if t_tensor >= t_hbm else 'hbm',
                          'frac_of_bound': round(max(t_tensor, t_hbm) * B / (ms_per_step * 1e-3), 4) if ms_per_step > 0 else None,
                          'launches_per_step': calls / a.steps})
        if l.index != 0:
            igemm_flops += fl * B
            igemm_ms += ms_per_step
    n_igemm = sum(int(p['launches_per_step']) for p in per_layer[1:])
    achieved = igemm_flops / (igemm_ms * 1e-3) / 1e12 if igemm_ms > 0 else 0.0
    peak = peaks['tflops_sustained']
    # DRAM bytes per launch from the committed ncu --set full capture of the same launches (profiles/, made by
    # tools/gpu_ncu_traffic.sh); only quoted when the capture's precision and arena chunk are this run's
    traffic = traffic_src = None
    tpath = os.path.join(ROOT, 'profiles', f'r01_igemm_traffic_chunk{a.chunk}.json')
    if os.path.exists(tpath):
        with open(tpath) as f:
            tj = json.load(f)
        if tj.get('precision') == a.precision and tj.get('chunk_images') == min(a.chunk, B):
            traffic = tj['dram_bytes_per_launch_avg']
            traffic_src = os.path.relpath(tpath, ROOT)
    roofline = {'bound': 'tensor', 'kernel': 'itmo::conv_igemm_kernel', 'achieved': round(achieved, 2),
                'peak': peak, 'unit': 'TFLOP/s', 'frac': round(achieved / peak, 4), 'traffic': traffic,
                'traffic_unit': 'DRAM bytes per launch (ncu dram__bytes_read.sum + dram__bytes_write.sum)',
                'traffic_source': traffic_src,
                'peak_source': f"{peaks['source']} (bf16 sustained; kernel timed inside a long step)",
                'flops_per_launch': igemm_flops / max(1, n_igemm),
                'avg_launch_ms': round(igemm_ms / max(1, n_igemm), 4),
                'share_of_step': round(igemm_ms / ms_step, 4),
                # every layer against ITS bound (tensor: bf16 sustained peak; hbm: measured copy bandwidth)
                'layers': [{k: p[k] for k in ('layer', 'ms_per_step', 'tflops', 'gbs', 'bound', 'frac_of_bound')}
                           for p in per_layer]}

    if rank == 0:
        log('per-layer (rank 0, CUDA events inside the timed region):')
        for p in per_layer:
            log(f"  {p['layer']:<22} {p['ms_per_step']:9.3f} ms/step  {p['tflops']:8.1f} TFLOP/s  {p['gbs']:8.1f} GB/s  "
                f"{p['bound']:<6} {p['frac_of_bound']}")
        log(f'  sum of layers {sum(p["ms_per_step"] for p in per_layer):.3f} ms, step {ms_step:.3f} ms')
        os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
        with open(os.path.join(ROOT, 'gpurun_out', f'bench_layers_{a.precision}_n{world}.json'), 'w') as f:
            json.dump({'batch': B, 'chunk': a.chunk, 'precision': a.precision, 'ms_per_step': ms_step,
                       'layers': per_layer}, f, indent=1)

    cpu_baseline = None
    if rank == 0 and world == 1 and not a.no_cpu:
        ips, n, cores, secs = cpu_reference_time(ws, a.cpu_sample, budget_s=a.cpu_budget)
        cpu_baseline = {'value': round(ips, 4), 'unit': UNIT, 'cores': cores, 'kind': 'port',
                        'sample': f'{n} of the {B} images, one image per predict call (predict.py:50-62), torch-CPU '
                                  f'fp32 restatement of model.py:204-278, {secs:.1f} s; TensorFlow-CPU unavailable '
                                  'on this box'}

    if rank == 0:
        line = {
            'metric': METRIC, 'value': round(value, 2), 'unit': UNIT, 'n_gpus': world, 'steps': a.steps,
            'warmup': max(a.warmup, 3) if a.warmup else 0, 'ms_per_step': round(ms_step, 3),
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': a.precision, 'data': 'synthetic',
            'mpix_per_s': round(value * MPIX_PER_IMAGE, 2),
            'tflops_per_gpu': round(value / world * graph.forward_flops_per_image() / 1e12, 2),
            'config': workload_config(a, a.chunk),
            'clocks': clocks,
            'e2e': {'value': round(e2e_value, 2), 'unit': UNIT, 'h2d_bytes_per_step': int(hx.nbytes),
                    'd2h_bytes_per_step': int(hy.nbytes), 'steps': e2e_steps,
                    'api': 'Model.predict_u8 -> itmo_predict_host (pinned host buffers)'},
            'gpu_launches': int(launches),
            'roofline': roofline,
            'cpu_baseline': cpu_baseline,
        }
        print(json.dumps(line), flush=True)
    model.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def run_frames_arm(a):
    """--workload 4k (BASELINE configs[3]): 3840x2160 uint8 frames, tiled 512x512 with a 96-pixel halo (84 tiles per
    frame), frames round-robin over the ranks, fused inverse-Reinhard output (float32 HDR). Reported in Mpix/s of
    UNIQUE output pixels (8.29 Mpix per frame); executed FLOPs are 84 x 165.37 GFLOP per frame."""
    import numpy as np
    import torch
    import torch.distributed as dist
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device')
    a.gpus = world
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)
    import __graft_entry__ as G
    if not os.path.exists(G.LIB):
        G.build()
    import cnn_itmo_b200 as M
    from cnn_itmo_b200 import frames
    from cnn_itmo_b200.model import pinned_empty
    H, W, F = 2160, 3840, a.frames
    model = M.Model(M.initial_weights(seed=218), precision=a.precision, device=local_rank, max_batch=min(a.chunk, 84))
    g = torch.Generator(device=dev).manual_seed(99 + rank)
    x = torch.randint(0, 256, (F, H, W, 3), dtype=torch.uint8, device=dev, generator=g)
    y = torch.empty((F, H, W, 3), dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream()

    def step():
        frames.predict_frames_device(model, x.data_ptr(), F, H, W, y.data_ptr(), 'hdr', 0.18, stream.cuda_stream)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    for _ in range(max(a.warmup, 3)):
        step()
    barrier()
    launches0 = model.launch_count()
    model.set_profiling(True)          # CUDA events around every layer launch, on the launch stream
    model.layer_times(reset=True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for _ in range(a.steps):
        step()
    ev1.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_total = ev0.elapsed_time(ev1)
    launches = model.launch_count() - launches0
    lt = model.layer_times(reset=True)
    model.set_profiling(False)
    # roofline of the implicit-GEMM kernel over the EXECUTED tiles (84 per frame), as in the batch workload
    from cnn_itmo_b200 import graph as _g
    ig_ms = sum(lt[l.index][0] for l in _g.LAYERS if l.kind != 'head' and l.index != 0)
    ig_calls = sum(lt[l.index][1] for l in _g.LAYERS if l.kind != 'head' and l.index != 0)
    ig_flops = (_g.forward_flops_per_image() - 2.0 * IMG * IMG * 27 * 32) * 84 * F * a.steps
    peaks = measured_peaks()
    ach = ig_flops / (ig_ms * 1e-3) / 1e12 if ig_ms > 0 else None
    roofline = {'bound': 'tensor', 'kernel': 'itmo::conv_igemm_kernel (executed tiles)', 'achieved': round(ach, 2) if ach else None,
                'peak': peaks['tflops_sustained'], 'unit': 'TFLOP/s',
                'frac': round(ach / peaks['tflops_sustained'], 4) if ach else None, 'traffic': None,
                'peak_source': peaks['source'] + ' (bf16 sustained)',
                'flops_per_launch': ig_flops / max(1, ig_calls), 'avg_launch_ms': round(ig_ms / max(1, ig_calls), 4),
                'share_of_step': round(ig_ms / ms_total, 4) if ms_total > 0 else None}
    if world > 1:
        t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    mpix = world * F * a.steps * (H * W / 1e6) / (ms_total / 1e3)
    # end to end: pinned host uint8 frames -> device -> HDR float32 -> pinned host
    hx = pinned_empty((F, H, W, 3), np.uint8)
    hy = pinned_empty((F, H, W, 3), np.float32)
    hx[...] = x.cpu().numpy()
    hxt, hyt = torch.from_numpy(hx), torch.from_numpy(hy)
    barrier()
    t0 = time.perf_counter()
    for _ in range(max(1, min(a.steps, 3))):
        x.copy_(hxt, non_blocking=True)
        step()
        hyt.copy_(y, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    e2e = world * F * max(1, min(a.steps, 3)) * (H * W / 1e6) / dt
    if rank == 0:
        from cnn_itmo_b200 import graph
        line = {'metric': '4K SDR->HDR frames, unique output pixels', 'value': round(mpix, 2), 'unit': 'Mpix/s',
                'n_gpus': world, 'steps': a.steps, 'warmup': max(a.warmup, 3), 'ms_per_step': round(ms_total / a.steps, 3),
                'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': a.precision,
                'data': 'synthetic',
                'frames_per_s': round(world * F * a.steps / (ms_total / 1e3), 3),
                'executed_tflops_per_gpu': round(F * a.steps * 84 * graph.forward_flops_per_image() / (ms_total / 1e3) / 1e12, 2),
                'config': {'workload': f'{F} synthetic 3840x2160x3 uint8 frames per GPU, 84 tiles of 512x512 per frame '
                                       '(96-pixel halo), fused inverse-Reinhard float32 HDR output (BASELINE configs[3])',
                           'precision': a.precision, 'parallelism': f'dp{world} (frames round-robin, no collective)',
                           'l2': 'each frame streams > 13 GB of activations; no explicit flush'},
                'clocks': clocks,
                'e2e': {'value': round(e2e, 2), 'unit': 'Mpix/s', 'h2d_bytes_per_step': int(hx.nbytes),
                        'd2h_bytes_per_step': int(hy.nbytes), 'api': 'frames.predict_frames_device with pinned host copies'},
                'gpu_launches': int(launches), 'roofline': roofline, 'cpu_baseline': None}
        print(json.dumps(line), flush=True)
    model.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def run_train_arm(a):
    """--workload train (BASELINE configs[4]): data-parallel training step on synthetic SDR/HDR pairs, GLOBAL batch
    a.train_batch (512) split over the ranks (strong scaling), micro-batched through the activation arena; forward in
    training mode + dgrad/wgrad + ONE NCCL all-reduce of the flat fp32 gradient buffer (44.6 MB) + fused RMSprop.
    `value` = images/s of whole optimizer steps with the batch resident in HBM; `e2e` = Trainer.train_on_batch on pinned
    host arrays (H2D of the batch and the loss read-back inside the timed region)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device')
    a.gpus = world
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)
    import __graft_entry__ as G
    if not os.path.exists(G.LIB):
        G.build()
    import cnn_itmo_b200 as M
    from cnn_itmo_b200 import graph, train
    from cnn_itmo_b200.model import pinned_empty
    if a.train_batch % world:
        raise SystemExit(f'bench.py: global batch {a.train_batch} is not divisible by {world} ranks')
    per_rank = a.train_batch // world
    mb = min(a.micro_batch, per_rank)
    tr = train.Trainer(M.initial_weights(seed=218), micro_batch=mb, image_hw=(IMG, IMG), device=local_rank,
                       optimizer='rmsprop')
    g = torch.Generator(device=dev).manual_seed(7 + rank)
    x = torch.randint(0, 256, (per_rank, IMG, IMG, 3), dtype=torch.uint8, device=dev, generator=g)
    y = torch.rand((per_rank, IMG, IMG, 3), dtype=torch.float32, device=dev, generator=g)
    loss_count = float(a.train_batch) * IMG * IMG * 3
    stream = torch.cuda.current_stream()
    state = {'step': 0}

    def step():
        tr.zero_grad()
        for i0 in range(0, per_rank, mb):
            tr.backward(x[i0:i0 + mb], y[i0:i0 + mb], loss_count, seed=state['step'] * 7919 + i0)
        tr.allreduce_grads()
        tr.update()
        state['step'] += 1

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    for _ in range(max(a.warmup, 3)):
        step()
    barrier()
    tr.metrics(reset=True)
    tr.set_profiling(True)
    tr.phase_times(reset=True)
    launches0 = tr.launch_count()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for _ in range(a.steps):
        step()
    ev1.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_total = ev0.elapsed_time(ev1)
    phases = tr.phase_times(reset=True)
    tr.set_profiling(False)
    launches = tr.launch_count() - launches0
    sq, hits = tr.metrics(reset=True)
    if world > 1:
        t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ips = a.train_batch * a.steps / (ms_total / 1e3)
    # end to end through the public call: pinned host batch -> H2D -> step -> loss back on the host
    hx = pinned_empty((per_rank, IMG, IMG, 3), np.uint8)
    hy = pinned_empty((per_rank, IMG, IMG, 3), np.float32)
    hx[...] = x.cpu().numpy()
    hy[...] = y.cpu().numpy()
    e2e_steps = max(1, min(a.steps, 3))
    barrier()
    t0 = time.perf_counter()
    last = None
    for _ in range(e2e_steps):
        last = tr.train_on_batch(hx, hy)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    e2e = a.train_batch * e2e_steps / dt
    if rank == 0:
        peaks = measured_peaks()
        fwd = graph.forward_flops_per_image()
        convs = [l for l in graph.LAYERS if l.kind != 'head']
        flop = lambda l: 2.0 * (l.hin ** 2) * (9 if l.kind == 'conv' else 4) * sum(l.cin) * l.cout
        f_fwd = sum(flop(l) for l in convs)
        f_dgrad = sum(flop(l) for l in convs[1:])          # the first layer needs no data gradient
        f_wgrad = f_fwd
        gemm_ms = phases['fwd_conv'][0] + phases['dgrad'][0] + phases['wgrad'][0]
        gemm_calls = phases['fwd_conv'][1] + phases['dgrad'][1] + phases['wgrad'][1]
        images_rank = per_rank * a.steps
        ach = (f_fwd + f_dgrad + f_wgrad) * images_rank / (gemm_ms / 1e3) / 1e12 if gemm_ms > 0 else None
        per_phase = {k: {'ms_per_step': round(v[0] / a.steps, 3), 'launches_per_step': v[1] // a.steps}
                     for k, v in phases.items()}
        per_phase['fwd_conv']['tflops'] = round(f_fwd * images_rank / (phases['fwd_conv'][0] / 1e3) / 1e12, 1)
        per_phase['dgrad']['tflops'] = round(f_dgrad * images_rank / (phases['dgrad'][0] / 1e3) / 1e12, 1)
        per_phase['wgrad']['tflops'] = round(f_wgrad * images_rank / (phases['wgrad'][0] / 1e3) / 1e12, 1)
        line = {'metric': 'U-Net SDR->HDR training throughput (512x512 images, forward + backward + RMSprop)',
                'value': round(ips, 2), 'unit': 'images/s', 'n_gpus': world, 'steps': a.steps,
                'warmup': max(a.warmup, 3), 'ms_per_step': round(ms_total / a.steps, 3), 'higher_is_better': True,
                'scaling': 'strong', 'vs_baseline': None, 'dtype': 'bf16', 'data': 'synthetic',
                'mpix_per_s': round(ips * MPIX_PER_IMAGE, 2),
                'tflops_per_gpu': round((f_fwd + f_dgrad + f_wgrad) * ips / world / 1e12, 1),
                'config': {'workload': f'data-parallel training step, global batch {a.train_batch} ({per_rank} per GPU in '
                                       f'micro-batches of {mb}), 512x512x3 uint8 SDR in, float32 target, random-init '
                                       'weights, Dropout on, RMSprop (BASELINE configs[4])',
                           'global_batch': a.train_batch, 'micro_batch': mb,
                           'parallelism': f'dp{world} (one NCCL all-reduce of the 44.6 MB flat fp32 gradient per step)',
                           'l2': 'every micro-batch streams > 10 GB of activations; no explicit flush'},
                'clocks': clocks,
                'e2e': {'value': round(e2e, 2), 'unit': 'images/s', 'h2d_bytes_per_step': int(hx.nbytes + hy.nbytes),
                        'd2h_bytes_per_step': 16, 'steps': e2e_steps, 'api': 'Trainer.train_on_batch (pinned host arrays)',
                        'last_loss': round(float(last[0]), 6)},
                'gpu_launches': int(launches),
                'roofline': {'bound': 'tensor', 'kernel': 'conv_igemm_kernel (forward + dgrad) + wgrad_igemm_kernel',
                             'achieved': round(ach, 1) if ach else None, 'peak': peaks['tflops_sustained'],
                             'unit': 'TFLOP/s', 'frac': round(ach / peaks['tflops_sustained'], 4) if ach else None,
                             'traffic': None, 'peak_source': peaks['source'] + ' (bf16 sustained)',
                             'flops_per_launch': (f_fwd + f_dgrad + f_wgrad) * images_rank / max(gemm_calls, 1),
                             'avg_launch_ms': round(gemm_ms / max(gemm_calls, 1), 4),
                             'share_of_step': round(gemm_ms / ms_total, 4), 'phases': per_phase},
                'train_loss': round(sq / (a.steps * per_rank * IMG * IMG * 3.0), 6), 'cpu_baseline': None}
        print(json.dumps(line), flush=True)
    tr.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--batch', type=int, default=256, help='images per GPU per step (BASELINE configs[2]: 256)')
    ap.add_argument('--chunk', type=int, default=int(os.environ.get('ITMO_BENCH_CHUNK', '64')),
                    help='images the activation arena holds (itmo_create max_batch)')
    ap.add_argument('--precision', default='bf16', choices=['bf16', 'fp16', 'tf32', 'fp32'])
    ap.add_argument('--cpu-sample', type=int, default=48,
                    help='upper bound on the images of the cpu_baseline sample (the --cpu-budget seconds decide); the '
                         'reference arm runs min(this, 12) images per step')
    ap.add_argument('--cpu-budget', type=float, default=15.0, help='seconds of CPU work for cpu_baseline')
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg')
    ap.add_argument('--train-batch', type=int, default=512, help='GLOBAL batch of --workload train (configs[4]: 512)')
    ap.add_argument('--micro-batch', type=int, default=32, help='images per forward/backward pass of --workload train')
    ap.add_argument('--workload', default='batch', choices=['batch', '4k', 'train'],
                    help="'batch' = BASELINE configs[2] (default, the headline); '4k' = configs[3] tiled frames")
    ap.add_argument('--frames', type=int, default=2, help='4K frames per GPU per step (--workload 4k)')
    a = ap.parse_args(argv)
    if a.impl == 'reference':
        return run_reference_arm(a)
    if a.workload == '4k':
        return run_frames_arm(a)
    if a.workload == 'train':
        return run_train_arm(a)
    return run_b200_arm(a)


if __name__ == '__main__':
    sys.exit(main())
