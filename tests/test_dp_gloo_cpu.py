"""World-size-2 gloo tests of the data-parallel host logic (phantom_vlb_b200/dp.py): sampler sharding, the gradient
all-reduce + 1/world folding contract of the optimiser front-end, the Pearson moment merge, and the voxel-sharded
readout head (feature all-gather, loss / d(hrf) / norm partial sums).  CPU only; the CUDA
kernels on either side of these collectives are covered by the -m gpu tests."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle.ref_model import pearson_from_stats, pearson_r, pearson_stats
from phantom_vlb_b200 import dp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert dp.world() == (rank, world)
        # ---- sampler: disjoint, equal-sized, identical permutation on every rank
        idx = dp.shard_indices(101, rank, world, seed=1234, epoch=3)
        gathered = [None] * world
        dist.all_gather_object(gathered, idx)
        flat = sum(gathered, [])
        assert len(set(flat)) == len(flat) == 100 and all(len(g) == 50 for g in gathered)
        # ---- gradient exchange: sum across ranks, then grad_scale = 1/world == mean over the global batch
        rng = np.random.RandomState(7)
        g_all = rng.randn(world, 1000).astype(np.float32)
        g = torch.from_numpy(g_all[rank].copy())
        w = dp.allreduce_sum_(g)
        assert w == world
        assert np.allclose(g.numpy() * (1.0 / w), g_all.mean(0), atol=1e-6)
        # global-norm clip computed AFTER the reduce is identical on all ranks -> no second collective
        norms = [None] * world
        dist.all_gather_object(norms, float((g * (1.0 / w)).norm()))
        assert norms[0] == norms[1]
        # ---- Pearson: row shards merged through the additive fp64 moments
        x, y = rng.randn(501, 33), rng.randn(501, 33)
        y[:, :9] += 0.5 * x[:, :9]
        x[5, 2] = np.nan
        rows = np.array_split(np.arange(501), world)[rank]            # ragged shards (251 / 250)
        st = torch.from_numpy(pearson_stats(x[rows], y[rows]))
        st, n = dp.merge_moments_(st, len(rows))
        assert n == 501
        r = pearson_from_stats(st.numpy(), n)
        assert np.allclose(r, pearson_r(x, y), atol=1e-12)
        # ---- config 4: readout head sharded by voxel. Features are all-gathered in rank order, each rank computes its
        # voxel slice for ALL clips, and only the loss partial sums, d(hrf) and the ridge-slice squared norm are summed.
        torch.manual_seed(3)
        Bl, H, P = 3, 16, 43
        feats_all = torch.randn(world * Bl, H, dtype=torch.float64)
        W = torch.randn(P, H, dtype=torch.float64, requires_grad=True)
        y_all = torch.randn(world * Bl, P, dtype=torch.float64)
        hrf_full = feats_all.clone().requires_grad_(True)
        loss_full = ((hrf_full @ W.t() - y_all) ** 2).mean() + 1e-3 * (W ** 2).sum()
        loss_full.backward()
        mine = feats_all[rank * Bl:(rank + 1) * Bl].contiguous()
        gathered = dp.gather_rows(mine)
        assert torch.equal(gathered, feats_all)
        reduce = dp.make_sum_reducer()
        assert reduce is not None
        lo, hi = dp.shard_columns(P, rank, world)
        Wl = W.detach()[lo:hi]
        d = gathered @ Wl.t() - y_all[:, lo:hi]
        acc = torch.tensor([float((d ** 2).sum()), float((Wl ** 2).sum())], dtype=torch.float64)
        reduce(acc)
        n_total = world * Bl * P
        assert abs(float(acc[0] / n_total + 1e-3 * acc[1]) - float(loss_full)) < 1e-12
        dpred = 2 * d / n_total
        dhrf = dpred @ Wl
        reduce(dhrf)
        assert torch.allclose(dhrf, hrf_full.grad, atol=1e-12)
        gW = dpred.t() @ gathered + 2e-3 * Wl
        assert torch.allclose(gW, W.grad[lo:hi], atol=1e-12)
        sq = torch.tensor([float((gW ** 2).sum())], dtype=torch.float64)
        reduce(sq)
        assert abs(float(sq) - float((W.grad ** 2).sum())) < 1e-9
        # ---- the engine's gradient hook: slices of one flat buffer handed over out of order (head first, then layer groups from
        # the top down) are reduced as ONE all-reduce after the backward ("deferred"), or one by one where they become final ("inline")
        calls = []
        real = dist.all_reduce

        def counting(t, *a, **k):
            calls.append(t.numel())
            return real(t, *a, **k)

        dist.all_reduce = counting
        try:
            for mode, want_calls in (("deferred", [100]), ("inline", [40, 30, 30])):
                flat = torch.from_numpy(g_all[rank, :100].copy())
                red = dp.OverlappedGradReducer(mode=mode)
                calls.clear()
                for lo_, hi_ in ((60, 100), (30, 60), (0, 30)):
                    red(flat[lo_:hi_])
                if mode == "deferred":
                    assert calls == []                              # nothing moves before wait()
                red.wait()
                assert calls == want_calls, (mode, calls)
                assert np.allclose(flat.numpy(), g_all[:, :100].sum(0), atol=1e-6)
                red.wait()                                          # idempotent: nothing pending
                assert calls == want_calls
        finally:
            dist.all_reduce = real
        q.put((rank, "ok"))
    except Exception as e:                                           # pragma: no cover
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_dp_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res


def test_shard_helpers_single_process():
    assert dp.world() == (0, 1)
    a = dp.shard_indices(10, 0, 3, shuffle=False, drop_last=False)
    b = dp.shard_indices(10, 1, 3, shuffle=False, drop_last=False)
    c = dp.shard_indices(10, 2, 3, shuffle=False, drop_last=False)
    assert len(a) == len(b) == len(c) == 4 and set(a + b + c) == set(range(10))
    cover = []
    for r in range(8):
        lo, hi = dp.shard_columns(250003, r, 8)
        assert lo % 8 == 0
        cover += [(lo, hi)]
    assert cover[0][0] == 0 and cover[-1][1] == 250003 and all(cover[i][1] == cover[i + 1][0] for i in range(7))
    s = dp.ShardedSampler(20, rank=1, world_size=4, seed=1)
    assert len(list(s)) == len(s) == 5
