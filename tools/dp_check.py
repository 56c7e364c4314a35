"""Data-parallel training check on N GPUs (run under torchrun): the all-reduced flat gradient of N ranks must equal the
gradient one rank accumulates over the same shards as micro-batches — BIT FOR BIT at N = 2 (the kernels use no
floating-point atomics and every gradient tensor receives one addition per micro-batch, so rank0 + rank1 is the same
fp32 sum as micro-batch 0 + micro-batch 1; for N > 2 NCCL's reduction tree may associate differently, so the bar there
is 1e-6 relative) — the single-rank path must reproduce itself exactly run to run, and every rank must end the step
with identical parameters.  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dp_check.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist

rank, world, local = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
if world > 1:
    dist.init_process_group('nccl', device_id=dev)
import cnn_itmo_b200 as M
from cnn_itmo_b200 import train
from oracle import unet_train_ref as T   # synthetic pairs only

per = 2
x, y = T.synthetic_pair(per * world, 512, 512, seed=3)
ws = M.initial_weights(11)
tr = train.Trainer(ws, micro_batch=per, device=local)
b, e = train.shard_range(per * world, world, rank)
count = train.global_loss_count(per, 512, 512, world)
tr.zero_grad()
tr.backward(torch.from_numpy(x[b:e]).to(dev), torch.from_numpy(y[b:e]).to(dev), count, seed=0, dropout=False)
tr.allreduce_grads()
g_dp = tr.grads.clone()
ref0 = g_dp.clone()
if world > 1:
    dist.broadcast(ref0, 0)
same = float((g_dp - ref0).abs().max())
tr.update()
p = tr.params.clone()
p0 = p.clone()
if world > 1:
    dist.broadcast(p0, 0)
same_p = float((p - p0).abs().max())
# single-rank reference: the same shards as successive micro-batches
tr2 = train.Trainer(ws, micro_batch=per, device=local)
tr2.zero_grad()
for r in range(world):
    bb, ee = train.shard_range(per * world, world, r)
    tr2.backward(torch.from_numpy(x[bb:ee]).to(dev), torch.from_numpy(y[bb:ee]).to(dev), count, seed=0, dropout=False)
torch.cuda.synchronize()
g_ref = tr2.grads.clone()
rel = float((g_dp - g_ref).norm() / g_ref.norm())
exact = bool(torch.equal(g_dp, g_ref))
# run-to-run: the single-rank path must reproduce itself exactly (fixed-order reductions everywhere)
tr2.zero_grad()
for r in range(world):
    bb, ee = train.shard_range(per * world, world, r)
    tr2.backward(torch.from_numpy(x[bb:ee]).to(dev), torch.from_numpy(y[bb:ee]).to(dev), count, seed=0, dropout=False)
torch.cuda.synchronize()
noise = float((tr2.grads - g_ref).norm() / g_ref.norm())
# ---- fused exchange (itmo_exchange_step: reduce-scatter -> optimizer -> all-gather in ONE kernel over peer memory):
# two optimizer steps on every rank's shard; all ranks must hold bit-identical parameters, equal — bit for bit, at any N,
# because the kernel sums the shards in rank order — to ONE rank accumulating the same shards and taking plain steps.
fused = {}
for opt in ('rmsprop', 'adam'):
    trf = train.Trainer(ws, micro_batch=per, device=local, optimizer=opt, exchange='fused')
    trs = train.Trainer(ws, micro_batch=per, device=local, optimizer=opt)
    for step in range(2):
        trf.zero_grad()
        trf.backward(torch.from_numpy(x[b:e]).to(dev), torch.from_numpy(y[b:e]).to(dev), count, seed=step, dropout=False)
        trf.exchange_and_update()
        trs.zero_grad()
        for r in range(world):
            bb, ee = train.shard_range(per * world, world, r)
            trs.backward(torch.from_numpy(x[bb:ee]).to(dev), torch.from_numpy(y[bb:ee]).to(dev), count, seed=step, dropout=False)
        trs.update()
    torch.cuda.synchronize()
    pf = trf.params.clone()
    pf0 = pf.clone()
    if world > 1:
        dist.broadcast(pf0, 0)
    st_f = trf.get_optimizer_weights()      # gathers the owners' shards
    st_s = trs.get_optimizer_weights()
    fused[opt] = {'params_bit_identical_across_ranks': bool(torch.equal(pf, pf0)),
                  'params_bit_identical_to_single_rank_steps': bool(torch.equal(pf, trs.params)),
                  'optimizer_state_bit_identical_to_single_rank': all(np.array_equal(a_, b_) for a_, b_ in zip(st_f, st_s)),
                  'moved': float((pf - torch.from_numpy(train.flatten_layers(ws)).to(dev)).abs().max())}
    trf.close()
    trs.close()
out = {'world': world, 'rank': rank, 'fused_exchange': fused, 'grad_rel_vs_single_rank': rel, 'grad_bit_identical_to_single_rank': exact,
       'single_rank_run_to_run_rel': noise,
       'grad_max_abs_diff_vs_rank0': same,
       'param_max_abs_diff_vs_rank0': same_p, 'grad_norm': float(g_dp.norm())}
if world > 1:
    gathered = [None] * world
    dist.all_gather_object(gathered, out)
    out = gathered
if rank == 0:
    print(json.dumps(out))
    ok = all((o['grad_bit_identical_to_single_rank'] if world <= 2 else o['grad_rel_vs_single_rank'] <= 1e-6) and
             o['single_rank_run_to_run_rel'] == 0.0 and
             o['grad_max_abs_diff_vs_rank0'] == 0.0 and
             o['param_max_abs_diff_vs_rank0'] == 0.0 and
             all(f['params_bit_identical_across_ranks'] and f['params_bit_identical_to_single_rank_steps'] and
                 f['optimizer_state_bit_identical_to_single_rank'] and f['moved'] > 0 for f in o['fused_exchange'].values())
             for o in (out if isinstance(out, list) else [out]))
    print('dp_check', 'OK' if ok else 'FAILED')
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
