"""Multi-GPU plumbing of the hot path: one process per GPU, the pose batch sharded across ranks.

Every sample (pose, mesh handle, K, gt crop, loss points) is independent through steps 1-5, so the data path has
NO collective: each rank renders and scores its own contiguous slice of the batch.  Two things cross ranks:
  * the logged loss: the reference's `loss.mean()` over the global batch (loss.py:15-16,121) = all-reduce(SUM) of the
    per-rank per-sample sums divided by the global batch — `global_loss_mean`;
  * Q2 (rotation_representation.py:165-168): the pose update uses the intrinsics of the FIRST sample of the batch for
    every sample, so every rank must use the GLOBAL first sample's fx, fy — `first_sample_intrinsics` broadcasts
    them from rank 0.
The reference has no distributed code at all; the CNN's gradient all-reduce is stock DDP.
Works with any torch.distributed backend (NCCL over NVLink on the GPU box; gloo in the CPU tests).
"""
import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_bounds(global_batch, rank=None, world_size=None):
    """Contiguous split of the batch dimension; the first `global_batch % world` ranks take one extra sample."""
    if rank is None or world_size is None:
        rank, world_size = world()
    assert 0 <= rank < world_size and global_batch >= 0
    base, extra = divmod(global_batch, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard(batch, rank=None, world_size=None):
    """Slices every array-like / list of a batch tuple or dict along dim 0 (mesh_paths is a list, the rest tensors)."""
    def cut(v, lo, hi):
        return v[lo:hi]
    if isinstance(batch, dict):
        n = len(next(iter(batch.values())))
        lo, hi = shard_bounds(n, rank, world_size)
        return {k: cut(v, lo, hi) for k, v in batch.items()}
    n = len(batch[0])
    lo, hi = shard_bounds(n, rank, world_size)
    return type(batch)(cut(v, lo, hi) for v in batch)


def global_loss_mean(per_sample, global_batch=None):
    """Mean over the GLOBAL batch of a per-sample loss vector held shard-wise.  Differentiable w.r.t. the local
    shard: d(mean)/d(per_sample[i]) = 1 / global_batch, which is what DDP's gradient averaging of per-rank means
    gives only for equal shards — this is exact for ragged shards too."""
    local_sum = per_sample.sum()
    count = torch.tensor([float(per_sample.numel())], device=per_sample.device, dtype=torch.float64)
    rank, ws = world()
    if ws == 1:
        n = float(global_batch) if global_batch is not None else float(per_sample.numel())
        return local_sum / n
    tot = local_sum.detach().to(torch.float64).reshape(1).clone()
    buf = torch.cat([tot, count])
    dist.all_reduce(buf, op=dist.ReduceOp.SUM)
    n = float(global_batch) if global_batch is not None else float(buf[1].item())
    # value = global mean; gradient flows only through the local sum
    return (local_sum - local_sum.detach()) / n + (buf[0] / n).to(per_sample.dtype)


def _collective_device():
    if dist.is_available() and dist.is_initialized() and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def first_sample_intrinsics(Ks_local):
    """(fx, fy) of the global batch's first sample (Q2), broadcast from rank 0; Ks_local (b,3,3) may be empty
    on other ranks."""
    rank, ws = world()
    # NCCL has no CPU backend: the 2-value buffer is staged on this rank's GPU wherever Ks lives (the reference's
    # cam_mats stay on the CPU, Q5); gloo (CPU tests) keeps it on the host
    dev = _collective_device()
    v = torch.zeros(2, dtype=torch.float64, device=dev)
    if rank == 0:
        K0 = torch.as_tensor(Ks_local)[0]
        v[0], v[1] = float(K0[0, 0]), float(K0[1, 1])
    if ws > 1:
        dist.broadcast(v, src=0)
    return float(v[0]), float(v[1])
