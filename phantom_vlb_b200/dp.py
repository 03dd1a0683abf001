"""Data-parallel host logic (one process per GPU, torch.distributed; NCCL on the box, gloo in CPU tests).

The path has exactly two exchange steps:
  * LoRA/head gradients: ONE all_reduce(sum) of the flat fp32 gradient buffer per step; the 1/world mean is folded
    into vlb_adamw_step(grad_scale) and the global-norm clip runs after the reduce (identical on all ranks).
    Reference: Lightning DDPStrategy around VLBLitModule (config/experiment/VLB_vllama2_friends_lora.yaml:41-50).
  * Pearson r over a row-sharded validation season: all_reduce(sum) of the additive fp64 moments [5, P] and of the
    row count, then finalize on every rank (reference concatenates everything on one device, src/utils.py:101-106).
Frozen weights are replicated and never communicated; clips are independent units.
"""
import os

import torch
import torch.distributed as dist


def world():
    """(rank, world_size) of the default process group, (0, 1) when not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_indices(n, rank, world_size, seed=0, epoch=0, shuffle=True, drop_last=True):
    """DistributedSampler-equivalent: one permutation from the shared (seed, epoch), strided by rank. Every rank gets
    the same count (drop_last) or the tail is padded by wrapping (not drop_last), so step counts match across ranks."""
    if shuffle:
        g = torch.Generator().manual_seed(int(seed) * 1000003 + int(epoch))
        perm = torch.randperm(n, generator=g).tolist()
    else:
        perm = list(range(n))
    if drop_last:
        per = n // world_size
        perm = perm[:per * world_size]
    else:
        per = (n + world_size - 1) // world_size
        perm = perm + perm[:per * world_size - n]
    return perm[rank::world_size]


class ShardedSampler(torch.utils.data.Sampler):
    """torch Sampler over shard_indices (what VLBDataModule uses when world_size > 1 outside Lightning)."""

    def __init__(self, n, rank=None, world_size=None, seed=0, shuffle=True, drop_last=True):
        r, w = world()
        self.n, self.rank, self.world_size = n, r if rank is None else rank, w if world_size is None else world_size
        self.seed, self.shuffle, self.drop_last, self.epoch = seed, shuffle, drop_last, 0

    def set_epoch(self, epoch):
        self.epoch = epoch

    def __iter__(self):
        return iter(shard_indices(self.n, self.rank, self.world_size, self.seed, self.epoch, self.shuffle, self.drop_last))

    def __len__(self):
        return self.n // self.world_size if self.drop_last else (self.n + self.world_size - 1) // self.world_size


def allreduce_sum_(flat, group=None):
    """In-place sum of a flat buffer across ranks (no-op for a single process). Returns the world size so the caller
    can fold 1/world into the optimiser kernel instead of a separate scaling pass."""
    if not (dist.is_available() and dist.is_initialized()):
        return 1
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return dist.get_world_size(group)


def merge_moments_(stats, n_rows, group=None):
    """Sum the per-rank Pearson moments [5, P] (fp64) and row counts; returns (stats, total_rows). The raw moments are
    additive, so the merge is exact in any order up to fp64 rounding."""
    n = torch.tensor([float(n_rows)], device=stats.device, dtype=torch.float64)
    if dist.is_available() and dist.is_initialized():
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(n, op=dist.ReduceOp.SUM, group=group)
    return stats, int(round(float(n.item())))


def shard_columns(P, rank, world_size, align=8):
    """[lo, hi) voxel slice owned by `rank` when the readout head is sharded by voxel (config 4: P ~ 250k).
    Slices are `align`-aligned (TMA needs 16-byte rows) and cover [0, P) exactly."""
    per = (P + world_size - 1) // world_size
    per = (per + align - 1) // align * align
    lo = min(P, rank * per)
    hi = min(P, lo + per)
    return lo, hi


def gather_rows(t, group=None):
    """All-gather along dim 0 in rank order: [B, ...] on every rank -> [world*B, ...] (the pooled features of config 4:
    every voxel shard of the readout head sees every clip of the global batch)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return t
    w = dist.get_world_size(group)
    out = torch.empty((w * t.shape[0],) + tuple(t.shape[1:]), device=t.device, dtype=t.dtype)
    dist.all_gather_into_tensor(out, t.contiguous(), group=group)
    return out


def make_sum_reducer(group=None):
    """callable(tensor) -> in-place sum across ranks, or None for a single process. Installed as Engine.shard_reduce when
    the readout head is sharded by voxel: it is applied to the loss partial sums [2], to d(hrf) [B, H] and to the squared
    gradient norm of the rank-private ridge slice."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return None

    def reduce(t):
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        return t
    return reduce


class OverlappedGradReducer:
    """Engine.grad_ready target: all-reduces every finished slice of the flat gradient ('launch per-chunk all-reduces as
    backward finishes each group of layers'). `wait()` must be called before the optimiser kernels. Three schedules:
      "overlap"   on a side stream while the backward of the earlier layers runs. Hides the transfer, but every GEMM of this library is
                  a persistent kernel sized to ALL SMs with a static tile order: while NCCL's CTAs hold a few SMs the last GEMM CTAs
                  start only when the first ones have finished their whole share -- on one 8-GPU box the step-start -> backward-end
                  time went from 512 to 542 ms (profiles/r02an_*).
      "inline"    on the compute stream, right where the slice becomes final: nothing runs beside NCCL, the ranks meet once per chunk.
      "deferred"  (default) one all-reduce of the union of the slices inside wait(), i.e. after the whole backward."""

    def __init__(self, group=None, device=None, timing=False, mode=None):
        self.group, self.timing = group, timing
        self.mode = mode or os.environ.get("VLB_DP_REDUCE", "deferred")
        assert self.mode in ("overlap", "inline", "deferred"), self.mode
        self.stream = torch.cuda.Stream(device) if self.mode == "overlap" else None
        self.spans = []                     # (start, end) events of each all-reduce (timing=True)
        self.pending = []                   # deferred slices
        self.bytes = 0

    def _reduce(self, t, stream):
        if self.timing:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        if self.timing:
            e1.record(stream)
            self.spans.append((e0, e1))
        self.bytes += t.numel() * t.element_size()

    def __call__(self, flat_slice):
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(self.group) == 1:
            return
        if self.mode == "deferred":
            self.pending.append(flat_slice)
        elif self.mode == "inline":
            self._reduce(flat_slice, self._cur())
        else:
            ready = torch.cuda.Event()
            ready.record()                  # on the compute stream: everything that writes this slice has been enqueued
            with torch.cuda.stream(self.stream):
                self.stream.wait_event(ready)
                self._reduce(flat_slice, self.stream)

    def wait(self):
        if self.mode == "overlap":
            torch.cuda.current_stream().wait_stream(self.stream)
        elif self.pending:
            # the slices are pieces of one flat buffer: merge neighbours so that the usual case is a single all-reduce
            parts = sorted(self.pending, key=lambda t: t.data_ptr())
            self.pending = []
            runs = [[parts[0]]]
            for t in parts[1:]:
                last = runs[-1][-1]
                if last.data_ptr() + last.numel() * last.element_size() == t.data_ptr() and t.dtype == last.dtype:
                    runs[-1].append(t)
                else:
                    runs.append([t])
            for run in runs:
                n = sum(t.numel() for t in run)
                self._reduce(run[0].as_strided((n,), (1,)) if len(run) > 1 else run[0], self._cur())

    def _cur(self):
        return torch.cuda.current_stream() if self.timing else None      # only the timing events need the stream (gloo tests have none)

    def pop_ms(self):
        """Sum of the all-reduce durations recorded since the last call (synchronise first)."""
        ms = sum(e0.elapsed_time(e1) for e0, e1 in self.spans)
        self.spans = []
        return ms
