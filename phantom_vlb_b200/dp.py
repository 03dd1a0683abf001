"""Data-parallel host logic (one process per GPU, torch.distributed; NCCL on the box, gloo in CPU tests).

The path has exactly two exchange steps:
  * LoRA/head gradients: ONE all_reduce(sum) of the flat fp32 gradient buffer per step; the 1/world mean is folded
    into vlb_adamw_step(grad_scale) and the global-norm clip runs after the reduce (identical on all ranks).
    Reference: Lightning DDPStrategy around VLBLitModule (config/experiment/VLB_vllama2_friends_lora.yaml:41-50).
  * Pearson r over a row-sharded validation season: all_reduce(sum) of the additive fp64 moments [5, P] and of the
    row count, then finalize on every rank (reference concatenates everything on one device, src/utils.py:101-106).
Frozen weights are replicated and never communicated; clips are independent units.
"""
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
    """Engine.grad_ready target: all-reduces every finished slice of the flat gradient on a side stream while the backward of the
    earlier decoder layers is still running ('launch per-chunk all-reduces as backward finishes each group of
    layers'). `wait()` makes the compute stream wait for the last one; it must be called before the optimiser kernels."""

    def __init__(self, group=None, device=None, timing=False):
        self.group, self.timing = group, timing
        self.stream = torch.cuda.Stream(device)
        self.spans = []                     # (start, end) events of each all-reduce on the side stream (timing=True)
        self.bytes = 0

    def __call__(self, flat_slice):
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(self.group) == 1:
            return
        ready = torch.cuda.Event()
        ready.record()                      # on the compute stream: everything that writes this slice has been enqueued
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(ready)
            if self.timing:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(self.stream)
            dist.all_reduce(flat_slice, op=dist.ReduceOp.SUM, group=self.group)
            if self.timing:
                e1.record(self.stream)
                self.spans.append((e0, e1))
        self.bytes += flat_slice.numel() * flat_slice.element_size()

    def wait(self):
        torch.cuda.current_stream().wait_stream(self.stream)

    def pop_ms(self):
        """Sum of the all-reduce durations recorded since the last call (synchronise first)."""
        ms = sum(e0.elapsed_time(e1) for e0, e1 in self.spans)
        self.spans = []
        return ms
