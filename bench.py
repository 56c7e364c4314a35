#!/usr/bin/env python
"""bench.py — throughput of the CNN-ITMO U-Net hot path (predict.py:62, model.py:204-278) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's sm_100a path
    python bench.py --impl reference [--steps K] [--warmup W]      # CPU restatement of the reference path
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N   # one rank per GPU

Headline (BASELINE.json configs[2]): a "step" is one forward pass of the whole U-Net over one batch of 256 synthetic
512x512 SDR images per GPU, uint8 in -> uint8 out, in the library's default precision (fp16 storage / operands, fp32
accumulation in TMEM). The weights are the calibrated synthetic checkpoint (he-normal kernels, BatchNormalization
statistics as training would leave them, negative gammas included) written to a Keras-layout HDF5 file and loaded
through ``load_model`` like predict.py:24 does — the committed checkpoint the reference names is absent from the
reference tree (.MISSING_LARGE_BLOBS). Rank 0 prints ONE JSON line:

  value      images/s over all ranks, inputs already resident in HBM (itmo_forward on device buffers)
  e2e        images/s through Model.predict_u8 -> itmo_predict_host: pinned HOST uint8 in, HOST uint8 out,
             H2D and D2H copies inside the timed region
  parity     PSNR / max-abs of this run's context against the CPU oracle on 8 images (outside the timed region)
  roofline   the implicit-GEMM conv kernel (17 launches per arena chunk = 99.7 % of the FLOPs): algorithmic
             FLOPs / summed CUDA-event duration of those launches inside the timed region, against the
             measured bf16 tensor peak of MEASURED_PEAKS.json
  cpu_baseline  the torch-CPU (oneDNN) restatement of the reference graph (oracle/unet_ref.py) timed on this
             box's host cores on a bounded sample — TensorFlow itself cannot be installed here
  secondary  short legs run after the headline, at every N: "train" (configs[4]: data-parallel training step, global
             batch 512, strong scaling, the all-reduce timed separately), "frames_4k" (configs[3]: tiled 3840x2160
             frames, fused inverse-Reinhard) and "strict_fp32" (the 3xTF32 mode that meets max-abs <= 1e-3)

The GPU is warmed with >= 2.5 s of the same forward before the timed region so that the window runs at the sustained
(power-capped) clock the roofline denominator was measured at. L2 hygiene: every step streams > 40 GB of activations
through a 126 MB L2 (inputs alone are 201 MB), so no explicit flush is needed; ``config.l2`` says so.
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
def calibrated_checkpoint(hz, seed: int = 218) -> str:
    """Path of the Keras-layout HDF5 file holding the calibrated synthetic checkpoint; rank 0 writes it (fixture
    generation by oracle/unet_ref.py, outside every timed region), the other ranks wait for it."""
    path = os.path.join(tempfile.gettempdir(), f'itmo_calibrated-{seed}.hdf5')   # 45 MB: not under gpurun_out/
    if hz.rank == 0 and not os.path.exists(path):
        from oracle import unet_ref as R
        import cnn_itmo_b200 as M
        os.makedirs(os.path.dirname(path), exist_ok=True)
        tmp = path + f'.tmp{os.getpid()}'
        M.write_checkpoint(tmp, R.calibrated_weights(seed))
        os.replace(tmp, path)
    hz.barrier()
    return path


def parity_against_oracle(model, weights, n_images: int = 8):
    """The CHECKER leg: this context's predict.py flow on n_images synthetic SDR images against the CPU oracle
    (torch float32 restatement of model.py:204-278). Outside every timed region."""
    import numpy as np
    from oracle import unet_ref as R
    imgs = R.synthetic_images_u8(n_images, seed=5)
    x = R.preprocess_u8(imgs)
    ref = R.forward_torch(R.TorchWeights(weights), x)
    out = model.predict(x)
    u8 = model.predict_u8(imgs)
    ref_u8 = R.postprocess_u8(ref)
    d = np.abs(u8.astype(np.int32) - ref_u8.astype(np.int32))
    return {'weights': 'calibrated', 'n_images': int(n_images), 'precision': model.precision,
            'psnr_db': round(R.psnr(out, ref), 2), 'max_abs': float(np.abs(out - ref).max()),
            'u8_psnr_db': round(R.psnr(u8 / 255.0, ref_u8 / 255.0), 2), 'u8_max_lsb': int(d.max()),
            'u8_frac_differing': round(float((d != 0).mean()), 5),
            'saturated_store_groups': int(model.saturation_count()),
            'oracle': 'oracle/unet_ref.py forward_torch (fp32, oneDNN); numerics unpinned: no TF/Keras, no checkpoint'}


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




def workload_config(a):
    """The SAME dict in both arms (the driver compares them): what is computed, and how each arm samples it."""
    return {'workload': f'full U-Net forward (19 weight layers, 165.37 GFLOP/image), batch {a.batch} per GPU, '
                        f'512x512x3 uint8 in -> uint8 out, calibrated synthetic checkpoint read from a Keras-layout '
                        f'HDF5 file (BASELINE configs[2])',
            'batch_per_gpu': a.batch, 'global_batch': a.batch * a.gpus,
            'parallelism': f'dp{a.gpus} (independent image batches, no collective on the data path)',
            'l2': 'inputs and activations far larger than L2 (>40 GB streamed per step); no explicit flush',
            'arena_chunk_images': a.chunk,
            'arms': {'b200': 'whole batch every step; storage / operand type in `dtype`, fp32 accumulation',
                     'reference': f'{REF_SAMPLE}-image sample of the batch per step, one image per predict call '
                                  '(predict.py:50-62), torch-CPU f32 restatement on all host threads'}}


REF_SAMPLE = 12   # 20 default steps x 12 images ~ 1 minute of CPU work


def run_reference_arm(a):
    """--impl reference: the reference's own CPU path cannot run here (TensorFlow 1.x / Keras 2.2 are not
    installable), so this times its CPU restatement (oracle/unet_ref.py, torch-CPU/oneDNN fp32) with all host
    threads. Each step = a bounded 12-image sample of the batch-256 workload, one image per predict call."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return 0
    import numpy as np
    import torch
    from oracle import unet_ref as R
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ws = R.calibrated_weights(218)
    tw = R.TorchWeights(ws)
    imgs = np.random.default_rng(0).integers(0, 256, (REF_SAMPLE, IMG, IMG, 3), dtype=np.uint8)
    for _ in range(max(1, a.warmup)):
        R.predict_like_reference(tw, imgs[:1])
    t0 = time.perf_counter()
    for _ in range(a.steps):
        R.predict_like_reference(tw, imgs)
    dt = time.perf_counter() - t0
    ips = a.steps * REF_SAMPLE / dt
    sample = (f'{REF_SAMPLE}-image sample of the {a.batch} images per step, one image per predict call '
              f'(predict.py:50-62), torch-CPU fp32 restatement, {cores} threads')
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': round(ips, 4), 'unit': UNIT, 'n_gpus': a.gpus,
        'steps': a.steps, 'warmup': a.warmup, 'ms_per_step': round(1e3 * dt / a.steps, 3), 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'mpix_per_s': round(ips * MPIX_PER_IMAGE, 4),
        'config': workload_config(a),
        'cpu_baseline': {'value': round(ips, 4), 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': round(ips, 4), 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
        'note': 'TensorFlow-CPU (the reference runtime) is unavailable on this box; this is its CPU restatement',
    }
    print(json.dumps(line), flush=True)
    return 0


# --------------------------------------------------------------------------------------------------
# process group / timing plumbing shared by the legs
# --------------------------------------------------------------------------------------------------
class Harness:
    """One rank of the job: device, process group (NCCL when world > 1), barrier and max-over-ranks."""

    def __init__(self, a):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get('WORLD_SIZE', '1'))
        self.rank = int(os.environ.get('RANK', '0'))
        self.local_rank = int(os.environ.get('LOCAL_RANK', '0'))
        if not torch.cuda.is_available():
            raise SystemExit('bench.py: no CUDA device — the sm_100a library is the only implementation '
                             '(use --impl reference for the CPU restatement)')
        if self.world != a.gpus:
            if self.world == 1 and a.gpus > 1:
                raise SystemExit(f'bench.py --gpus {a.gpus}: launch with torchrun --nproc-per-node {a.gpus} '
                                 '(one rank per GPU)')
            a.gpus = self.world
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device('cuda', self.local_rank)
        if self.world > 1:
            pin_rank_to_numa_cpus(self.local_rank)
        if self.world > 1:
            os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
            dist.init_process_group('nccl', device_id=self.dev)
        import __graft_entry__ as G
        if not os.path.exists(G.LIB):
            G.build()
        self.stream = torch.cuda.current_stream()

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()

    def max_over_ranks(self, v: float) -> float:
        if self.world == 1:
            return float(v)
        t = self.torch.tensor([v], device=self.dev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, step, steps: int, sampler=None):
        """EXACTLY `steps` calls of step() between two CUDA events on the launch stream, barrier + synchronize on both
        sides; returns the max over ranks of the elapsed milliseconds."""
        torch = self.torch
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        if sampler is not None:
            sampler.start()
        ev0.record(self.stream)
        for _ in range(steps):
            step()
        ev1.record(self.stream)
        self.barrier()
        clocks = sampler.stop() if sampler is not None else None
        return self.max_over_ranks(ev0.elapsed_time(ev1)), clocks

    def close(self):
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()


def pin_rank_to_numa_cpus(local_rank: int):
    """Keep this rank's host threads (pinned-memory copies, launches) on the CPUs nearest its GPU when the topology
    says so; a no-op when the box reports one NUMA node for every GPU (as this pool's B200 boxes do)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        n = os.cpu_count() or 1
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, (n + 63) // 64)
        cpus = [i for i in range(n) if (mask[i // 64] >> (i % 64)) & 1]
        if cpus and len(cpus) < n:
            os.sched_setaffinity(0, cpus)
    except Exception:  # noqa: BLE001
        pass


def warm_up(step, sync, min_steps: int, min_seconds: float) -> int:
    """`min_steps` untimed steps, then more until `min_seconds` of device work have passed: the timed window then runs
    at the sustained (power-capped) clock, not at the boost clock of an idle board."""
    n = 0
    t0 = time.perf_counter()
    while n < min_steps or time.perf_counter() - t0 < min_seconds:
        step()
        n += 1
        if n % 4 == 0 or n >= min_steps:
            sync()
    sync()
    return n


# --------------------------------------------------------------------------------------------------
# headline leg: BASELINE configs[2]
# --------------------------------------------------------------------------------------------------
def layer_table(graph, layer_times, steps, B, precision, peaks):
    """Every layer against ITS bound from the CUDA events of the timed region."""
    per_layer = []
    igemm_flops = igemm_ms = 0.0
    # conv2d_1 + conv2d_2 run as ONE kernel (csrc/stem_fused.cuh) for uint8 input in the 16-bit modes: no launch is
    # booked under conv2d_1 then, and conv2d_2's row carries both layers' FLOPs and the pair's HBM bytes
    stem_fused = layer_times[0][1] == 0 and layer_times[1][1] > 0
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
        if stem_fused and l.index == 0:
            per_layer.append({'layer': l.name, 'ms_per_step': 0.0, 'fused_into': 'conv2d_2 (stem_fused_kernel)'})
            continue
        if stem_fused and l.index == 1:
            fl += 2.0 * IMG * IMG * 27 * 32
        ms_per_step = ms / steps
        tf = fl * B / (ms_per_step * 1e-3) / 1e12 if ms_per_step > 0 else 0.0
        # algorithmic HBM bytes per image: every source read once, the output written once (+ the
        # pooled copy), the weights once per launch; uint8 pixels in front of conv2d_1, uint8 RGB behind conv2d_14
        es = 2 if precision in ('bf16', 'fp16') else 4
        hout = 2 * l.hin if l.kind == 'convT' else l.hin
        by = l.hin ** 2 * (3 if (l.index == 0 or (stem_fused and l.index == 1)) else sum(l.cin) * es)
        by += IMG * IMG * 3 if l.index == 17 else hout ** 2 * l.cout * es * (1.25 if l.pool else 1.0)
        gbs = by * B / (ms_per_step * 1e-3) / 1e9 if ms_per_step > 0 else 0.0
        t_tensor = fl / (peaks['tflops_sustained'] * 1e12)
        t_hbm = by / (peaks['hbm_gbs'] * 1e9)
        per_layer.append({'layer': l.name, 'ms_per_step': round(ms_per_step, 4), 'tflops': round(tf, 1),
                          'gbs': round(gbs, 1), 'bound': 'tensor' if t_tensor >= t_hbm else 'hbm',
                          'frac_of_bound': round(max(t_tensor, t_hbm) * B / (ms_per_step * 1e-3), 4) if ms_per_step > 0 else None,
                          'launches_per_step': calls / steps})
        if l.index != 0:    # conv2d_1 alone is rgb_conv_tc_kernel, not the implicit-GEMM kernel family
            igemm_flops += fl * B
            igemm_ms += ms_per_step
    return per_layer, igemm_flops, igemm_ms


def run_batch_leg(a, hz, precision=None, batch=None, steps=None, with_e2e=True, with_parity=True, warm_s=None):
    """Full U-Net forward over `batch` images per GPU. Returns the JSON line (rank 0) or None."""
    import numpy as np
    torch = hz.torch
    import cnn_itmo_b200 as M
    from cnn_itmo_b200 import graph
    from cnn_itmo_b200.model import pinned_empty

    peaks = measured_peaks()
    B = batch or a.batch
    steps = steps or a.steps
    ckpt = calibrated_checkpoint(hz)
    # the public load path of predict.py:24; precision None = the library default (fp16)
    model = M.load_model(ckpt, precision=precision or a.precision, device=hz.local_rank, max_batch=min(a.chunk, B))
    prec = model.precision
    g = torch.Generator(device=hz.dev).manual_seed(1234 + hz.rank)
    x = torch.randint(0, 256, (B, IMG, IMG, 3), dtype=torch.uint8, device=hz.dev, generator=g)
    y = torch.empty_like(x)

    def step():
        model.forward_device(x.data_ptr(), 'u8', y.data_ptr(), 'u8', B, stream=hz.stream.cuda_stream)

    n_warm = warm_up(step, torch.cuda.synchronize, max(a.warmup, 3), a.warm_seconds if warm_s is None else warm_s)
    hz.barrier()
    model.set_profiling(True)
    model.layer_times(reset=True)
    launches0 = model.launch_count()
    ms_total, clocks = hz.timed(step, steps, ClockSampler(hz.local_rank) if hz.rank == 0 else None)
    launches = model.launch_count() - launches0
    layer_times = model.layer_times(reset=True)
    model.set_profiling(False)
    ms_step = ms_total / steps
    value = hz.world * B * steps / (ms_total / 1e3)

    probe = y[0, ::64, ::64].float()
    if not bool(torch.isfinite(probe).all()):
        raise SystemExit('bench.py: non-finite output')
    saturated = model.saturation_count()

    # ---- end to end: pinned host uint8 -> predict_u8 -> pinned host uint8 --------------------------
    e2e = None
    if with_e2e:
        hx = pinned_empty((B, IMG, IMG, 3), np.uint8)
        hy = pinned_empty((B, IMG, IMG, 3), np.uint8)
        hx[...] = x.cpu().numpy()
        model.predict_u8(hx, out=hy)  # warm-up (allocates the device staging buffers)
        e2e_steps = max(1, min(steps, 5))
        hz.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            model.predict_u8(hx, out=hy)
        torch.cuda.synchronize()
        dt = hz.max_over_ranks(time.perf_counter() - t0)
        same = bool((torch.from_numpy(np.asarray(hy[:2])).to(hz.dev) == y[:2]).all())
        if not same and not os.environ.get('ITMO_DBG'):  # ITMO_DBG timing experiments compute garbage on purpose
            raise SystemExit('bench.py: host path and device path disagree')
        e2e = {'value': round(hz.world * B * e2e_steps / dt, 2), 'unit': UNIT, 'h2d_bytes_per_step': int(hx.nbytes),
               'd2h_bytes_per_step': int(hy.nbytes), 'steps': e2e_steps,
               'api': 'Model.predict_u8 -> itmo_predict_host (pinned host buffers)'}
        del hx, hy

    per_layer, igemm_flops, igemm_ms = layer_table(graph, layer_times, steps, B, prec, peaks)
    n_igemm = sum(int(p.get('launches_per_step', 0)) for p in per_layer[1:])
    achieved = igemm_flops / (igemm_ms * 1e-3) / 1e12 if igemm_ms > 0 else 0.0
    peak = peaks['tflops_sustained']
    # DRAM bytes per launch from the committed ncu --set full capture of the same launches (profiles/, made by
    # tools/gpu_ncu_traffic.sh); only quoted when the capture's precision class and arena chunk are this run's
    traffic = traffic_src = None
    for rnd in ('r02', 'r01'):
        tpath = os.path.join(ROOT, 'profiles', f'{rnd}_igemm_traffic_chunk{a.chunk}.json')
        if os.path.exists(tpath):
            with open(tpath) as f:
                tj = json.load(f)
            same_bytes = {tj.get('precision'), prec} <= {'bf16', 'fp16'} or tj.get('precision') == prec
            if same_bytes and tj.get('chunk_images') == min(a.chunk, B):
                traffic = tj['dram_bytes_per_launch_avg']
                traffic_src = os.path.relpath(tpath, ROOT) + f" (captured in {tj.get('precision')}: same bytes)"
                break
    roofline = {'bound': 'tensor', 'kernel': 'itmo::conv_igemm_kernel (+ itmo::stem_fused_kernel for the first two layers, itmo::conv_rows_kernel for conv2d_3 / conv2d_4 / conv2d_14)', 'achieved': round(achieved, 2),
                'peak': peak, 'unit': 'TFLOP/s', 'frac': round(achieved / peak, 4), 'traffic': traffic,
                'traffic_unit': 'DRAM bytes per launch (ncu dram__bytes_read.sum + dram__bytes_write.sum)',
                'traffic_source': traffic_src,
                'peak_source': f"{peaks['source']} (bf16/fp16 sustained; kernel timed inside a long step after "
                               f"{n_warm} warm-up steps)",
                'flops_per_launch': igemm_flops / max(1, n_igemm),
                'avg_launch_ms': round(igemm_ms / max(1, n_igemm), 4),
                'share_of_step': round(igemm_ms / ms_step, 4),
                'layers': [{k: p[k] for k in ('layer', 'ms_per_step', 'tflops', 'gbs', 'bound', 'frac_of_bound', 'fused_into')
                            if k in p} for p in per_layer]}

    line = None
    if hz.rank == 0:
        log(f'per-layer [{prec}, batch {B}] (rank 0, CUDA events inside the timed region):')
        for p in per_layer:
            if 'fused_into' in p:
                log(f"  {p['layer']:<22} fused into {p['fused_into']}")
                continue
            log(f"  {p['layer']:<22} {p['ms_per_step']:9.3f} ms/step  {p['tflops']:8.1f} TFLOP/s  {p['gbs']:8.1f} GB/s  "
                f"{p['bound']:<6} {p['frac_of_bound']}")
        log(f'  sum of layers {sum(p["ms_per_step"] for p in per_layer):.3f} ms, step {ms_step:.3f} ms')
        os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
        with open(os.path.join(ROOT, 'gpurun_out', f'bench_layers_{prec}_n{hz.world}.json'), 'w') as f:
            json.dump({'batch': B, 'chunk': a.chunk, 'precision': prec, 'ms_per_step': ms_step,
                       'layers': per_layer}, f, indent=1)
        parity = None
        if with_parity:
            parity = parity_against_oracle(model, model._weights, a.parity_images)
            log(f'parity [{prec}]: {parity}')
        line = {
            'metric': METRIC, 'value': round(value, 2), 'unit': UNIT, 'n_gpus': hz.world, 'steps': steps,
            'warmup': a.warmup, 'warmup_steps_run': n_warm, 'ms_per_step': round(ms_step, 3),
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': prec, 'data': 'synthetic',
            'mpix_per_s': round(value * MPIX_PER_IMAGE, 2),
            'tflops_per_gpu': round(value / hz.world * graph.forward_flops_per_image() / 1e12, 2),
            'config': workload_config(a),
            'clocks': clocks,
            'e2e': e2e,
            'parity': parity,
            'saturated_store_groups': int(saturated),
            'gpu_launches': int(launches),
            'roofline': roofline,
        }
    ws = model._weights
    model.close()
    del x, y
    torch.cuda.empty_cache()
    return line, ws


# --------------------------------------------------------------------------------------------------
# BASELINE configs[3]: 4K frames
# --------------------------------------------------------------------------------------------------
def run_frames_leg(a, hz, weights=None, steps=None, warm_s=1.0):
    """3840x2160 uint8 frames, tiled 512x512 with a 96-pixel halo (84 tiles per frame), frames sharded over the ranks
    (no collective), fused inverse-Reinhard output (float32 HDR, like the floats inverse_Reinhard.m
    writes; the half-float output kind is timed beside it under e2e.half_float_output). Reported in Mpix/s of UNIQUE output pixels (8.29 Mpix per frame); executed
    FLOPs are 84 x 165.37 GFLOP per frame. Input frames are virtual-camera renders of seeded HDR fields when the
    synthesis kernels are available (datagen), else uniform noise."""
    import numpy as np
    torch = hz.torch
    import cnn_itmo_b200 as M
    from cnn_itmo_b200 import frames, graph
    from cnn_itmo_b200.model import pinned_empty
    steps = steps or a.steps
    H, W, F = 2160, 3840, a.frames
    if weights is None:
        weights = M.read_checkpoint(calibrated_checkpoint(hz))[0]
    model = M.Model(weights, precision=a.precision, device=hz.local_rank, max_batch=min(a.chunk, 84))
    prec = model.precision
    x, data = synth_frames(hz, F, H, W, seed=99 + hz.rank)
    y = torch.empty((F, H, W, 3), dtype=torch.float32, device=hz.dev)

    def step():
        frames.predict_frames_device(model, x.data_ptr(), F, H, W, y.data_ptr(), 'hdr', 0.18, hz.stream.cuda_stream)

    n_warm = warm_up(step, torch.cuda.synchronize, max(min(a.warmup, 3), 1), warm_s)
    hz.barrier()
    launches0 = model.launch_count()
    model.set_profiling(True)          # CUDA events around every layer launch, on the launch stream
    model.layer_times(reset=True)
    ms_total, clocks = hz.timed(step, steps, ClockSampler(hz.local_rank) if hz.rank == 0 else None)
    launches = model.launch_count() - launches0
    lt = model.layer_times(reset=True)
    model.set_profiling(False)
    if not bool(torch.isfinite(y[0, ::97, ::89]).all()):
        raise SystemExit('bench.py: non-finite HDR output')
    # roofline of the implicit-GEMM kernel over the EXECUTED tiles (84 per frame), as in the batch workload
    ig_ms = sum(lt[l.index][0] for l in graph.LAYERS if l.kind != 'head' and l.index != 0)
    ig_calls = sum(lt[l.index][1] for l in graph.LAYERS if l.kind != 'head' and l.index != 0)
    ig_flops = (graph.forward_flops_per_image() - 2.0 * IMG * IMG * 27 * 32) * 84 * F * steps
    peaks = measured_peaks()
    ach = ig_flops / (ig_ms * 1e-3) / 1e12 if ig_ms > 0 else None
    roofline = {'bound': 'tensor', 'kernel': 'itmo::conv_igemm_kernel (executed tiles)', 'achieved': round(ach, 2) if ach else None,
                'peak': peaks['tflops_sustained'], 'unit': 'TFLOP/s',
                'frac': round(ach / peaks['tflops_sustained'], 4) if ach else None, 'traffic': None,
                'peak_source': peaks['source'] + ' (bf16/fp16 sustained)',
                'flops_per_launch': ig_flops / max(1, ig_calls), 'avg_launch_ms': round(ig_ms / max(1, ig_calls), 4),
                'share_of_step': round(ig_ms / ms_total, 4) if ms_total > 0 else None}
    mpix = hz.world * F * steps * (H * W / 1e6) / (ms_total / 1e3)
    # end to end: pinned host uint8 frames -> device -> HDR float32 -> pinned host
    hx = pinned_empty((F, H, W, 3), np.uint8)
    hx[...] = x.cpu().numpy()
    # the public streaming call: every step's frames go host -> device and its HDR frames device -> host inside the timed
    # region; the upload of step k + 1 and the download of step k - 1 run under the kernels of step k (three streams)
    e2e_steps = max(2 * steps, 6)

    def stream_of(kind):
        n_out = 0
        for res in frames.predict_frame_batches(model, (hx for _ in range(e2e_steps)), out=kind):
            n_out += res.shape[0]
        assert n_out == F * e2e_steps
    stream_of('hdr')                   # warm-up: streams, device and pinned buffers of the call
    torch.cuda.synchronize()
    hz.barrier()
    t0 = time.perf_counter()
    stream_of('hdr')
    torch.cuda.synchronize()
    dt = hz.max_over_ranks(time.perf_counter() - t0)
    e2e = hz.world * F * e2e_steps * (H * W / 1e6) / dt
    # the same with the half-float output kind (OpenEXR's HALF pixel type): half the device -> host bytes, which is what
    # bounds this path when eight ranks share one host's PCIe / pinned-memory path
    stream_of('hdr_f16')
    torch.cuda.synchronize()
    hz.barrier()
    t0 = time.perf_counter()
    stream_of('hdr_f16')
    torch.cuda.synchronize()
    dth = hz.max_over_ranks(time.perf_counter() - t0)
    e2e_half = hz.world * F * e2e_steps * (H * W / 1e6) / dth
    line = None
    if hz.rank == 0:
        line = {'metric': '4K SDR->HDR frames, unique output pixels', 'value': round(mpix, 2), 'unit': 'Mpix/s',
                'n_gpus': hz.world, 'steps': steps, 'warmup': a.warmup, 'warmup_steps_run': n_warm,
                'ms_per_step': round(ms_total / steps, 3),
                'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': prec,
                'data': data,
                'frames_per_s': round(hz.world * F * steps / (ms_total / 1e3), 3),
                'executed_tflops_per_gpu': round(F * steps * 84 * graph.forward_flops_per_image() / (ms_total / 1e3) / 1e12, 2),
                'config': {'workload': f'{F} synthetic 3840x2160x3 uint8 frames per GPU, 84 tiles of 512x512 per frame '
                                       '(96-pixel halo), fused inverse-Reinhard float32 HDR output (BASELINE configs[3])',
                           'parallelism': f'dp{hz.world} (frames round-robin, no collective)',
                           'l2': 'each frame streams > 13 GB of activations; no explicit flush'},
                'clocks': clocks,
                'e2e': {'value': round(e2e, 2), 'unit': 'Mpix/s', 'h2d_bytes_per_step': int(hx.nbytes),
                        'd2h_bytes_per_step': int(hx.nbytes) * 4, 'steps': e2e_steps,
                        'api': 'frames.predict_frame_batches (pinned host batches; upload / kernels / download on three streams)',
                        'half_float_output': {'value': round(e2e_half, 2), 'unit': 'Mpix/s',
                                              'd2h_bytes_per_step': int(hx.nbytes) * 2, 'out_kind': 'hdr_f16'}},
                'gpu_launches': int(launches), 'roofline': roofline, 'cpu_baseline': None}
    model.close()
    del x, y, hx
    torch.cuda.empty_cache()
    return line


def synth_frames(hz, n, h, w, seed):
    """uint8 SDR frames on the device: virtual-camera renders (m-files/virtual_camera.m:13-35) of seeded HDR fields
    made by the library's synthesis kernels; falls back to uniform noise while those are not built."""
    torch = hz.torch
    try:
        from cnn_itmo_b200 import datagen
        sdr, _ = datagen.synthetic_pairs_device(n, h, w, seed=seed, device=hz.local_rank, want_target=False)
        return sdr, 'synthetic (virtual-camera renders of seeded HDR fields, datagen kernels)'
    except ImportError:
        g = torch.Generator(device=hz.dev).manual_seed(seed)
        return (torch.randint(0, 256, (n, h, w, 3), dtype=torch.uint8, device=hz.dev, generator=g),
                'synthetic (uniform noise)')


# --------------------------------------------------------------------------------------------------
# BASELINE configs[4]: data-parallel training step
# --------------------------------------------------------------------------------------------------
def synth_pairs(hz, n, seed):
    """(uint8 SDR input, float32 Reinhard target) on the device: the library's synthesis kernels
    (virtual_camera.m:13-35 / Reinhard.m:12-26 on seeded HDR fields); uniform noise while they are not built."""
    torch = hz.torch
    try:
        from cnn_itmo_b200 import datagen
        x, y = datagen.synthetic_pairs_device(n, IMG, IMG, seed=seed, device=hz.local_rank, want_target=True)
        return x, y, 'synthetic SDR/HDR pairs (virtual camera + Reinhard target kernels on seeded HDR fields)'
    except ImportError:
        g = torch.Generator(device=hz.dev).manual_seed(seed)
        x = torch.randint(0, 256, (n, IMG, IMG, 3), dtype=torch.uint8, device=hz.dev, generator=g)
        y = torch.rand((n, IMG, IMG, 3), dtype=torch.float32, device=hz.dev, generator=g)
        return x, y, 'synthetic (uniform noise)'


def run_train_leg(a, hz, steps=None, warmup=None):
    """Data-parallel training step on synthetic SDR/HDR pairs, GLOBAL batch a.train_batch (512) split over the ranks
    (strong scaling), micro-batched through the activation arena; forward in training mode + dgrad/wgrad + ONE NCCL
    all-reduce of the flat fp32 gradient buffer (44.6 MB) + fused RMSprop. `value` = images/s of whole optimizer steps
    with the batch resident in HBM; `e2e` = Trainer.train_on_batch on pinned host arrays (H2D of the batch and the loss
    read-back inside the timed region). The all-reduce is bracketed by its own CUDA events."""
    import numpy as np
    torch = hz.torch
    import cnn_itmo_b200 as M
    from cnn_itmo_b200 import graph, train
    from cnn_itmo_b200.model import pinned_empty
    steps = steps or a.steps
    warmup = max(a.warmup, 3) if warmup is None else warmup
    world = hz.world
    if a.train_batch % world:
        raise SystemExit(f'bench.py: global batch {a.train_batch} is not divisible by {world} ranks')
    per_rank = a.train_batch // world
    mb = min(a.micro_batch, per_rank)
    # gradient exchange: the fused peer-memory kernel (reduce-scatter -> RMSprop on the local shard -> all-gather, one
    # launch, no NCCL on the data path) when the ranks can map each other's windows, else NCCL all-reduce + optimizer
    tr = None
    exchange = a.exchange
    if exchange in ('auto', 'fused'):
        try:
            tr = train.Trainer(M.initial_weights(seed=218), micro_batch=mb, image_hw=(IMG, IMG), device=hz.local_rank,
                               optimizer='rmsprop', seed=0, exchange='fused')
            exchange = 'fused'
        except Exception as e:  # noqa: BLE001  (raised on every rank together, see Trainer.__init__)
            if a.exchange == 'fused':
                raise
            log(f'fused exchange unavailable ({e!r}); using NCCL all-reduce')
            exchange = 'nccl'
    if tr is None:
        tr = train.Trainer(M.initial_weights(seed=218), micro_batch=mb, image_hw=(IMG, IMG), device=hz.local_rank,
                           optimizer='rmsprop', seed=0)
    x, y, data = synth_pairs(hz, per_rank, seed=7 + hz.rank)
    loss_count = float(a.train_batch) * IMG * IMG * 3
    comm_events = []
    starts = list(range(0, per_rank, mb))

    def step():
        s = tr.step_index
        tr.zero_grad()
        for k, i0 in enumerate(starts):
            tr.backward(x[i0:i0 + mb], y[i0:i0 + mb], loss_count, seed=tr.micro_seed(s, i0),
                        freeze_moving=(k + 1 < len(starts)))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(hz.stream)
        if exchange == 'fused':
            tr.exchange_and_update()       # ONE kernel: reduce-scatter + RMSprop + all-gather (+ the operand repack)
            e1.record(hz.stream)
        else:
            tr.allreduce_grads()
            e1.record(hz.stream)
            tr.update()
        comm_events.append((e0, e1))

    for _ in range(warmup):
        step()
    hz.barrier()
    comm_events.clear()
    tr.metrics(reset=True)
    tr.set_profiling(True)
    tr.phase_times(reset=True)
    launches0 = tr.launch_count()
    ms_total, clocks = hz.timed(step, steps, ClockSampler(hz.local_rank) if hz.rank == 0 else None)
    phases = tr.phase_times(reset=True)
    tr.set_profiling(False)
    launches = tr.launch_count() - launches0
    sq, hits = tr.metrics(reset=True)
    allreduce_ms = hz.max_over_ranks(sum(e0.elapsed_time(e1) for e0, e1 in comm_events) / max(1, len(comm_events)))
    ips = a.train_batch * steps / (ms_total / 1e3)
    # end to end through the public call: pinned host batch -> H2D (overlapped with the previous micro-batch) -> step
    # -> loss back on the host
    hx = pinned_empty((per_rank, IMG, IMG, 3), np.uint8)
    hy = pinned_empty((per_rank, IMG, IMG, 3), np.float32)
    hx[...] = x.cpu().numpy()
    hy[...] = y.cpu().numpy()
    # H2D alone, for the record
    hz.barrier()
    t0 = time.perf_counter()
    x.copy_(torch.from_numpy(hx), non_blocking=True)
    y.copy_(torch.from_numpy(hy), non_blocking=True)
    torch.cuda.synchronize()
    h2d_ms = hz.max_over_ranks((time.perf_counter() - t0) * 1e3)
    e2e_steps = max(1, min(steps, 3))
    tr.train_on_batch(hx, hy, prefetch=(hx, hy))   # warm-up of the host path (side stream, allocator); the loop below is then in
                                                   # the steady state of a fit_generator epoch (every step finds its first micro-batch staged)
    hz.barrier()
    t0 = time.perf_counter()
    last = None
    for i in range(e2e_steps):
        # like Model.fit_generator's look-ahead: the next step's first micro-batch is uploaded under this step's compute
        # (every step still copies its whole batch host -> device inside the timed region)
        last = tr.train_on_batch(hx, hy, prefetch=(hx, hy) if i + 1 < e2e_steps else None)
    torch.cuda.synchronize()
    dt = hz.max_over_ranks(time.perf_counter() - t0)
    e2e = a.train_batch * e2e_steps / dt
    line = None
    if hz.rank == 0:
        peaks = measured_peaks()
        convs = [l for l in graph.LAYERS if l.kind != 'head']
        flop = lambda l: 2.0 * (l.hin ** 2) * (9 if l.kind == 'conv' else 4) * sum(l.cin) * l.cout
        f_fwd = sum(flop(l) for l in convs)
        f_dgrad = sum(flop(l) for l in convs[1:])          # the first layer needs no data gradient
        f_wgrad = f_fwd
        gemm_ms = phases['fwd_conv'][0] + phases['dgrad'][0] + phases['wgrad'][0]
        gemm_calls = phases['fwd_conv'][1] + phases['dgrad'][1] + phases['wgrad'][1]
        images_rank = per_rank * steps
        ach = (f_fwd + f_dgrad + f_wgrad) * images_rank / (gemm_ms / 1e3) / 1e12 if gemm_ms > 0 else None
        per_phase = {k: {'ms_per_step': round(v[0] / steps, 3), 'launches_per_step': v[1] // steps}
                     for k, v in phases.items()}
        for k, f in (('fwd_conv', f_fwd), ('dgrad', f_dgrad), ('wgrad', f_wgrad)):
            if phases[k][0] > 0:
                per_phase[k]['tflops'] = round(f * images_rank / (phases[k][0] / 1e3) / 1e12, 1)
        whole = (f_fwd + f_dgrad + f_wgrad) * a.train_batch / world / (ms_total / steps / 1e3) / 1e12
        line = {'metric': 'U-Net SDR->HDR training throughput (512x512 images, forward + backward + RMSprop)',
                'value': round(ips, 2), 'unit': 'images/s', 'n_gpus': world, 'steps': steps,
                'warmup': warmup, 'ms_per_step': round(ms_total / steps, 3), 'higher_is_better': True,
                'scaling': 'strong', 'vs_baseline': None, 'dtype': 'bf16', 'data': data,
                'mpix_per_s': round(ips * MPIX_PER_IMAGE, 2),
                'tflops_per_gpu': round(whole, 1),
                'whole_step_frac_of_peak': round(whole / peaks['tflops_sustained'], 4),
                'allreduce_ms': round(allreduce_ms, 3), 'h2d_ms': round(h2d_ms, 3),
                'exchange': {'mode': exchange,
                             'what_allreduce_ms_times': ('itmo_exchange_step: reduce-scatter + optimizer + all-gather over '
                                                         'NVLink peer memory + operand repack (includes waiting for the '
                                                         'slowest rank)') if exchange == 'fused' else
                             'the NCCL all-reduce alone (includes waiting for the slowest rank)'},
                'config': {'workload': f'data-parallel training step, global batch {a.train_batch} ({per_rank} per GPU in '
                                       f'micro-batches of {mb}), 512x512x3 uint8 SDR in, float32 target, random-init '
                                       'weights, Dropout on, RMSprop (BASELINE configs[4])',
                           'global_batch': a.train_batch, 'micro_batch': mb,
                           'parallelism': f'dp{world} (' + ('fused reduce-scatter / RMSprop / all-gather kernel over NVLink peer '
                                                             'memory' if exchange == 'fused' else
                                                             'one NCCL all-reduce of the 44.6 MB flat fp32 gradient per step') + ')',
                           'l2': 'every micro-batch streams > 10 GB of activations; no explicit flush'},
                'clocks': clocks,
                'e2e': {'value': round(e2e, 2), 'unit': 'images/s', 'h2d_bytes_per_step': int(hx.nbytes + hy.nbytes),
                        'd2h_bytes_per_step': 16, 'steps': e2e_steps, 'api': 'Trainer.train_on_batch (pinned host arrays; next batch prefetched like Model.fit_generator)',
                        'last_loss': round(float(last[0]), 6)},
                'gpu_launches': int(launches),
                'roofline': {'bound': 'tensor', 'kernel': 'conv_igemm_kernel (forward + dgrad) + wgrad_igemm_kernel',
                             'achieved': round(ach, 1) if ach else None, 'peak': peaks['tflops_sustained'],
                             'unit': 'TFLOP/s', 'frac': round(ach / peaks['tflops_sustained'], 4) if ach else None,
                             'traffic': None, 'peak_source': peaks['source'] + ' (bf16 sustained)',
                             'flops_per_launch': (f_fwd + f_dgrad + f_wgrad) * images_rank / max(gemm_calls, 1),
                             'avg_launch_ms': round(gemm_ms / max(gemm_calls, 1), 4),
                             'share_of_step': round(gemm_ms / ms_total, 4), 'phases': per_phase},
                'train_loss': round(sq / (steps * per_rank * IMG * IMG * 3.0), 6), 'cpu_baseline': None}
    tr.close()
    del x, y, hx, hy
    torch.cuda.empty_cache()
    return line


def compact(line, keys=('value', 'unit', 'ms_per_step', 'steps', 'scaling', 'dtype', 'data', 'e2e', 'roofline',
                        'allreduce_ms', 'exchange', 'h2d_ms', 'tflops_per_gpu', 'whole_step_frac_of_peak', 'frames_per_s',
                        'executed_tflops_per_gpu', 'gpu_launches', 'train_loss', 'parity', 'config')):
    if line is None:
        return None
    out = {k: line[k] for k in keys if k in line}
    if isinstance(out.get('roofline'), dict):
        out['roofline'] = {k: v for k, v in out['roofline'].items() if k != 'layers'}
    if isinstance(out.get('config'), dict):
        out['config'] = {'workload': out['config'].get('workload')}
    return out


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--batch', type=int, default=256, help='images per GPU per step (BASELINE configs[2]: 256)')
    ap.add_argument('--chunk', type=int, default=int(os.environ.get('ITMO_BENCH_CHUNK', '64')),
                    help='images the activation arena holds (itmo_create max_batch)')
    ap.add_argument('--precision', default=None, choices=['bf16', 'fp16', 'tf32', 'fp32'],
                    help='default: the library default (fp16)')
    ap.add_argument('--cpu-sample', type=int, default=48,
                    help='upper bound on the images of the cpu_baseline sample (the --cpu-budget seconds decide)')
    ap.add_argument('--cpu-budget', type=float, default=15.0, help='seconds of CPU work for cpu_baseline')
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg')
    ap.add_argument('--parity-images', type=int, default=8, help='images of the parity leg against the oracle')
    ap.add_argument('--train-batch', type=int, default=512, help='GLOBAL batch of the training leg (configs[4]: 512)')
    ap.add_argument('--exchange', default='auto', choices=['auto', 'fused', 'nccl'],
                    help='gradient exchange of the training leg: fused peer-memory kernel, NCCL all-reduce, or fused with '
                         'NCCL as the fallback when CUDA IPC is unavailable (auto)')
    ap.add_argument('--micro-batch', type=int, default=32, help='images per forward/backward pass of the training leg')
    ap.add_argument('--workload', default='batch', choices=['batch', '4k', 'train'],
                    help="'batch' = BASELINE configs[2] (default, the headline, followed by the secondary legs); "
                         "'4k' / 'train' = that leg alone, printed as the JSON line")
    ap.add_argument('--frames', type=int, default=2, help='4K frames per GPU per step')
    ap.add_argument('--no-secondary', action='store_true', help='headline leg only')
    ap.add_argument('--warm-seconds', type=float, default=2.5,
                    help='minimum duration of the untimed warm-up of the headline leg (0 for profiler runs)')
    ap.add_argument('--no-parity', action='store_true', help='skip the parity leg (profiler runs)')
    ap.add_argument('--secondary-steps', type=int, default=3, help='timed steps of each secondary leg')
    a = ap.parse_args(argv)
    if a.impl == 'reference':
        return run_reference_arm(a)
    hz = Harness(a)
    if a.workload == '4k':
        line = run_frames_leg(a, hz)
    elif a.workload == 'train':
        line = run_train_leg(a, hz)
    else:
        line, ws = run_batch_leg(a, hz, with_parity=not a.no_parity, with_e2e=not a.no_parity)
        secondary = {}
        if not a.no_secondary:
            t0 = time.perf_counter()
            for name, fn in (('train', lambda: run_train_leg(a, hz, steps=a.secondary_steps, warmup=2)),
                             ('frames_4k', lambda: run_frames_leg(a, hz, weights=ws, steps=a.secondary_steps)),
                             ('strict_fp32', lambda: run_batch_leg(a, hz, precision='fp32', batch=min(a.batch, 64),
                                                                   steps=a.secondary_steps, with_e2e=False,
                                                                   warm_s=0.5)[0])):
                try:
                    secondary[name] = compact(fn())
                except SystemExit:
                    raise
                except Exception as e:  # noqa: BLE001  (a secondary leg must not take the headline down)
                    log(f'secondary leg {name} failed: {e!r}')
                    secondary[name] = {'error': repr(e)}
                    hz.torch.cuda.empty_cache()
            log(f'secondary legs: {time.perf_counter() - t0:.1f} s')
        if hz.rank == 0:
            line['secondary'] = secondary
            line['cpu_baseline'] = None
            if hz.world == 1 and not a.no_cpu:
                ips, n, cores, secs = cpu_reference_time(ws, a.cpu_sample, budget_s=a.cpu_budget)
                line['cpu_baseline'] = {
                    'value': round(ips, 4), 'unit': UNIT, 'cores': cores, 'kind': 'port',
                    'sample': f'{n} of the {a.batch} images, one image per predict call (predict.py:50-62), torch-CPU '
                              f'fp32 restatement of model.py:204-278, {secs:.1f} s; TensorFlow-CPU unavailable on this box'}
    if hz.rank == 0 and line is not None:
        print(json.dumps(line), flush=True)
    hz.close()
    return 0


if __name__ == '__main__':
    sys.exit(main())
