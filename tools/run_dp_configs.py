"""BASELINE configs[3] / configs[4] on N GPUs of one box (one process per GPU, NCCL):

  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/run_dp_configs.py c4 [B] [P]
      frozen 7B backbone, 16-frame clips, B clips per rank -> pooled features -> all-gather -> readout head sharded by voxel
      (each rank owns P/N voxels), head-only AdamW step.  Also checks the sharded loss against an unsharded head.
  torchrun ... tools/run_dp_configs.py c5 [N_rows] [P]
      per-voxel Pearson r over a held-out season, rows sharded over the ranks, fp64 moments merged with one all-reduce.
Times are CUDA events, max over ranks. Rank 0 prints one JSON line per phase.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from bench import host_batches
from phantom_vlb_b200 import dp, ops


def setup():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    return rank, world, dev


def timed(fn, warm, it, world, dev):
    for _ in range(warm):
        fn()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        fn()
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / it], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms)


def measure_c4(rank, world, dev, B=4, P=250000, W=None, out=print):
    """BASELINE configs[3] on `world` ranks; `W` = an already resident frozen 7B replica (bench.py passes its own)."""
    from phantom_vlb_b200.engine import Engine
    from phantom_vlb_b200.model import ModelConfig, Trainables, frozen_random
    mc = ModelConfig(num_frames=16, num_target=P, use_lora=False)
    if W is None:
        W = frozen_random(mc, dev, 1234)                   # identical frozen replica on every rank
    lo, hi = dp.shard_columns(P, rank, world)
    T = Trainables(mc, dev, target_slice=(lo, hi))
    T.init_default(1234 + rank)
    eng = Engine(mc, W, T, device=dev)
    eng.shard_reduce = dp.make_sum_reducer()
    batch = {k: v.to(dev) for k, v in host_batches(mc, B, 1, seed=1234 + rank)[0].items()}     # rank-private clips
    u = torch.empty((B, mc.hidden), device=dev)
    sw = torch.empty((B,), device=dev)
    ms_f = timed(lambda: eng.extract_features(batch, out_u=u, out_sw=sw), 2, 4, world, dev)
    if rank == 0:
        out(json.dumps(dict(kind="c4_extract_features", n_gpus=world, clips_per_gpu=B, frames=16, ms=round(ms_f, 2),
                              clips_per_s=round(world * B / ms_f * 1e3, 2), tflops_per_gpu=round(39.0 * B / ms_f * 1e3, 1))), flush=True)
    g = torch.Generator(device=dev).manual_seed(99)        # same targets on every rank; each keeps its voxel columns
    y_all = torch.randn((world * B, P), device=dev, generator=g)
    y = y_all[:, lo:hi].contiguous()
    del y_all

    def head_step():
        ua, swa = dp.gather_rows(u), dp.gather_rows(sw)
        eng.cached_loss_and_grads(ua, swa, y)
        eng.optimizer_step(1e-4)

    ms_h = timed(head_step, 2, 5, world, dev)
    loss = float(eng.buf("hd_loss", (3,), torch.float32)[0])
    if rank == 0:
        per_rank = 34 * mc.hidden * (hi - lo)              # 6 B/param head traffic + 28 B/param AdamW
        out(json.dumps(dict(kind="c4_sharded_head_step", n_gpus=world, samples=world * B, P=P, voxels_per_gpu=hi - lo,
                              ms=round(ms_h, 3), loss=loss, alg_GB_per_gpu=round(per_rank / 1e9, 2),
                              achieved_GBs_per_gpu=round(per_rank / ms_h / 1e6, 1))), flush=True)
    # sharded == unsharded: same features, small head, rank 0 holds the full one
    Ps = 4096
    mcs = ModelConfig(num_frames=16, num_target=Ps, use_lora=False)
    los, his = dp.shard_columns(Ps, rank, world)
    gen = torch.Generator(device=dev).manual_seed(5)
    Wfull = torch.randn((Ps, mc.hidden), device=dev, generator=gen) * 0.02
    ys = torch.randn((world * B, Ps), device=dev, generator=gen)
    ua, swa = dp.gather_rows(u), dp.gather_rows(sw)

    def small(slice_, reducer):
        Ts = Trainables(mcs, dev, target_slice=slice_)
        Ts.init_default(1)
        Ts.p("ridge_w").copy_(Wfull[slice_[0]:slice_[1]]); Ts.p("ridge_b").zero_()
        Ts.sync_working()
        e = Engine(mcs, W, Ts, device=dev)
        e.shard_reduce = reducer
        out, _ = e.cached_loss_and_grads(ua, swa, ys[:, slice_[0]:slice_[1]].contiguous())
        return float(out[0]), Ts.g("ln2_w").clone()

    l_sh, g_sh = small((los, his), dp.make_sum_reducer())
    l_full, g_full = small((0, Ps), None)
    err = abs(l_sh - l_full) / abs(l_full)
    gerr = float((g_sh - g_full).norm() / g_full.norm())
    if rank == 0:
        out(json.dumps(dict(kind="c4_sharded_vs_unsharded", n_gpus=world, loss_rel_err=err, ln2_grad_rel_err=gerr)), flush=True)
    assert err < 1e-5 and gerr < 1e-3


def c4(B=4, P=250000):
    rank, world, dev = setup()
    measure_c4(rank, world, dev, B, P)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def measure_c5(rank, world, dev, N=22528, P=250000, out=print):
    """BASELINE configs[4]: rows of the held-out season sharded over the ranks, fp64 moments merged with one all-reduce."""
    rows = N // world + (1 if rank < N % world else 0)      # ragged row shards
    g = torch.Generator(device=dev).manual_seed(3 + rank)
    x = torch.randn((rows, P), device=dev, generator=g)
    y = 0.5 * x + torch.randn((rows, P), device=dev, generator=g)
    stats = torch.zeros((5, P), device=dev, dtype=torch.float64)
    res = {}

    def run():
        stats.zero_()
        ops.pearson(x, y, stats=stats, finalize=False)
        st, n = dp.merge_moments_(stats, rows)
        res["r"], res["n"] = ops.pearson_finalize(st, n), n

    ms = timed(run, 2, 5, world, dev)
    assert res["n"] == N
    r = res["r"]
    if rank == 0:
        nbytes = 2 * N * P * 4
        out(json.dumps(dict(kind="c5_pearson_row_sharded", n_gpus=world, N=N, P=P, ms=round(ms, 3), alg_GB=round(nbytes / 1e9, 2),
                              aggregate_GBs=round(nbytes / ms / 1e6, 1), mean_r=float(r.mean()), expect_r=round(0.5 / 1.25 ** 0.5, 4))), flush=True)
    assert abs(float(r.mean()) - 0.5 / 1.25 ** 0.5) < 5e-3


def c5(N=22528, P=250000):
    rank, world, dev = setup()
    measure_c5(rank, world, dev, N, P)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    which = sys.argv[1]
    args = [int(a) for a in sys.argv[2:]]
    {"c4": c4, "c5": c5}[which](*args)
