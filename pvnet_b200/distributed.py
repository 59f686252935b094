"""Multi-GPU voting: shard by image, no data-path collective, one packed gather at the end.

The reference's only parallel strategy is single-process torch.nn.DataParallel over the batch dimension
(tools/train_linemod.py:222-223): scatter images to GPUs, gather the [b,vn,2] keypoints on GPU 0.  Here it is one
process per GPU (torchrun), each rank votes on its own contiguous slice of the batch, and ONE collective
(all_gather_into_tensor over NCCL/NVLink; gloo in the CPU tests) assembles keypoints (+ covariances) on every rank.
Images are independent, so there is no exchange step inside the path; the draw streams are keyed by the GLOBAL image
id (pvv_job.image_id_base), which makes results independent of the number of ranks.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_bounds(global_b: int, world: int, rank: int):
    """Contiguous partition of [0, global_b): the first (global_b % world) ranks get one extra image."""
    base, rem = divmod(global_b, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_packed(local: torch.Tensor, global_b: int, group=None) -> torch.Tensor:
    """local [b_local, ...] float32 on every rank (contiguous partition by shard_bounds) -> [global_b, ...] on every rank.
    Uneven shards are padded to the largest shard so that a single all_gather_into_tensor is enough."""
    world = dist.get_world_size(group)
    if world == 1:
        return local
    per = -(-global_b // world)
    tail = tuple(local.shape[1:])
    send = local
    if local.shape[0] != per:
        send = torch.zeros((per,) + tail, dtype=local.dtype, device=local.device)
        send[: local.shape[0]] = local
    recv = torch.empty((world * per,) + tail, dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(recv, send.contiguous(), group=group)
    if global_b % world == 0:
        return recv
    parts = []
    for r in range(world):
        lo, hi = shard_bounds(global_b, world, r)
        parts.append(recv[r * per: r * per + (hi - lo)])
    return torch.cat(parts, 0)


class AsyncGather:
    """The packed gather of the results, taken OFF the voting stream.

    Round 1 issued `all_gather_into_tensor` on the compute stream after every step: ~30 us of pure collective latency
    (2.3 kB payload) between the finalisation of step i and the compaction of step i + 1, 7 % of a 0.4 ms step at 8 GPUs.
    Here the collective is launched with async_op=True -- NCCL's own stream waits (by event) for the voting stream to
    have produced `local`, the voting stream does not wait for NCCL -- into one of `depth` rotating receive buffers, so
    step i + 1 starts at once and overlaps the gather of step i.  `wait(slot)` makes the current stream wait for that
    gather (no host block) and returns the assembled tensor; a slot is waited before its buffer is reused.
    Equal shards only (global_b divisible by the world size): the padded path of gather_packed() stays synchronous."""

    def __init__(self, global_b, tail, device, dtype=torch.float32, group=None, depth=2):
        self.group, self.world = group, (dist.get_world_size(group) if dist.is_initialized() else 1)
        if global_b % self.world:
            raise ValueError("AsyncGather needs equal shards (global_b divisible by the world size)")
        self.bufs = [torch.empty((global_b,) + tuple(tail), dtype=dtype, device=device) for _ in range(depth)]
        self.work = [None] * depth
        self.n = 0

    def submit(self, local):
        slot = self.n % len(self.bufs)
        self.n += 1
        if self.work[slot] is not None:      # the buffer's previous gather must be complete before NCCL overwrites it
            self.work[slot].wait()
            self.work[slot] = None
        if self.world == 1:
            self.bufs[slot].copy_(local)
        else:
            self.work[slot] = dist.all_gather_into_tensor(self.bufs[slot], local.contiguous(), group=self.group, async_op=True)
        return slot

    def wait(self, slot):
        if self.work[slot] is not None:
            self.work[slot].wait()
            self.work[slot] = None
        return self.bufs[slot]

    def drain(self):
        for s in range(len(self.bufs)):
            self.wait(s)


def _shared_seed(seed, group, device):
    """seed=None: a fresh draw stream per call (like the reference's .random_()), drawn on rank 0 and broadcast so that
    every rank keys its draws with the same (seed, global image id)."""
    if seed is not None:
        return int(seed)
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return int(torch.randint(0, 2 ** 62, (1,), generator=_seed_gen()).item())
    t = torch.randint(0, 2 ** 62, (1,), generator=_seed_gen()).to(device)
    dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    return int(t.item())


_GEN = None


def _seed_gen():
    global _GEN
    if _GEN is None:
        _GEN = torch.Generator(device="cpu")
        _GEN.manual_seed(torch.initial_seed() & (2 ** 63 - 1))
    return _GEN


def ransac_voting_layer_v3_sharded(mask_local, vertex_local, round_hyp_num, global_b, *, group=None, seed=None,
                                   with_cov=False, cov_kwargs=None, vote_fn=None, cov_fn=None, gather=None, **kwargs):
    """Each rank passes ITS slice (shard_bounds) of the batch.  Returns keypoints [global_b,vn,2] (and, with
    with_cov=True, covariances [global_b,vn,2,2] from estimate_voting_distribution_with_mean) on every rank.
    `seed=None` draws a fresh seed on rank 0 and broadcasts it (one tiny collective); pass a seed to avoid it.
    `vote_fn` / `cov_fn` default to the CUDA entry points; the CPU tests inject stand-ins to exercise the plumbing.
    `gather`: an AsyncGather -- the call then returns the SLOT of the pending gather instead of tensors (the packed
    [global_b, vn, 2 or 6] result is `gather.wait(slot)`: keypoints [:, :, :2], covariances [:, :, 2:] row-major)."""
    if vote_fn is None or (with_cov and cov_fn is None):
        from . import ransac_voting_gpu as pv
        vote_fn = vote_fn or pv.ransac_voting_layer_v3
        cov_fn = cov_fn or pv.estimate_voting_distribution_with_mean
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    lo, hi = shard_bounds(global_b, world, rank)
    seed = _shared_seed(seed, group, mask_local.device)
    if mask_local.shape[0] != hi - lo:
        raise ValueError(f"rank {rank} owns images [{lo},{hi}) but was given {mask_local.shape[0]}")
    kp = vote_fn(mask_local, vertex_local, round_hyp_num, seed=seed, image_id_base=lo, **kwargs)
    if not with_cov:
        if gather is not None:
            return gather.submit(kp)
        return gather_packed(kp, global_b, group) if world > 1 else kp
    _, cov = cov_fn(mask_local, vertex_local, kp, seed=seed + 1, image_id_base=lo, **(cov_kwargs or {}))
    packed = torch.cat([kp, cov.reshape(cov.shape[0], cov.shape[1], 4)], 2)  # [b_local, vn, 6]: one collective
    if gather is not None:
        return gather.submit(packed)
    full = gather_packed(packed, global_b, group) if world > 1 else packed
    return full[:, :, :2].contiguous(), full[:, :, 2:].reshape(full.shape[0], full.shape[1], 2, 2)
