"""Data-parallel sharding of the synthesis hot path: one process per GPU, no collective on the data path.

The path partitions over utterances: every utterance's CFM decode and vocoding is independent, so
ranks take disjoint subsets of the batch and the only exchange is the final gather of mels / waveforms to one rank
(NCCL over NVLink on GPUs; the same code runs over gloo in the CPU tests).  The reference has no multi-GPU inference
path (matcha/cli.py synthesises on one device), so the interface here follows torch.distributed conventions.

    parts = partition(lengths, world)           # balanced by frames: cost of the path is ~linear in mel frames
    idx = parts[rank]
    wav_local = synthesise(mu[idx], mask[idx], ...)            # matcha_b200 hot path on this rank's GPU
    wav = gather_padded(wav_local, idx, len(lengths), dst=0)   # rank 0: (N, 1, L_max) in the caller's order
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence

import torch
import torch.distributed as dist


def partition(lengths: Sequence[int], world_size: int) -> List[List[int]]:
    """Longest-processing-time assignment of utterances to ranks, balanced by total frames.

    Deterministic (ties broken by index), every rank gets a (possibly empty) list, and within a rank the utterances
    stay sorted by decreasing length so that a rank can cut its shard into length-homogeneous micro-batches.
    """
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    order = sorted(range(len(lengths)), key=lambda i: (-int(lengths[i]), i))
    load = [0] * world_size
    parts: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], len(parts[k]), k))
        parts[r].append(i)
        load[r] += int(lengths[i])
    return parts


def micro_batches(idx: Sequence[int], lengths: Sequence[int], max_frames: int) -> List[List[int]]:
    """Cut one rank's shard (sorted by decreasing length) into batches whose padded size B * T_max <= max_frames."""
    out: List[List[int]] = []
    cur: List[int] = []
    t_max = 0
    for i in idx:
        t = max(t_max, int(lengths[i]))
        if cur and (len(cur) + 1) * t > max_frames:
            out.append(cur)
            cur, t = [], int(lengths[i])
        cur.append(i)
        t_max = t
    if cur:
        out.append(cur)
    return out


def gather_padded(local: torch.Tensor, idx: Sequence[int], n_total: int, dst: int = 0,
                  group: Optional[dist.ProcessGroup] = None) -> Optional[torch.Tensor]:
    """Gather per-rank outputs (b_local, C, L_local) to `dst` as (n_total, C, L_max) in the original utterance order.

    Shapes differ between ranks (shard sizes and padded lengths), so ranks first exchange (b_local, L_local) and the
    indices they hold, then send one padded buffer each.  Returns None on the other ranks.
    """
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if local.dim() != 3:
        raise ValueError("expected (batch, channels, length)")
    if local.shape[0] != len(idx):
        raise ValueError("one output row per index")
    if world == 1:
        out = local.new_zeros(n_total, local.shape[1], local.shape[2])
        out[torch.as_tensor(list(idx), dtype=torch.long, device=local.device)] = local
        return out
    dev = local.device
    meta = torch.tensor([local.shape[0], local.shape[2]], dtype=torch.long, device=dev)
    metas = [torch.empty_like(meta) for _ in range(world)]
    dist.all_gather(metas, meta, group=group)
    b_max = max(int(m[0]) for m in metas)
    l_max = max(int(m[1]) for m in metas)
    ids = torch.full((b_max,), -1, dtype=torch.long, device=dev)
    ids[: len(idx)] = torch.as_tensor(list(idx), dtype=torch.long, device=dev)
    buf = local.new_zeros(b_max, local.shape[1], l_max)
    buf[: local.shape[0], :, : local.shape[2]] = local
    id_list = [torch.empty_like(ids) for _ in range(world)] if rank == dst else None
    buf_list = [torch.empty_like(buf) for _ in range(world)] if rank == dst else None
    dist.gather(ids, id_list, dst=dst, group=group)
    dist.gather(buf, buf_list, dst=dst, group=group)
    if rank != dst:
        return None
    out = local.new_zeros(n_total, local.shape[1], l_max)
    seen = torch.zeros(n_total, dtype=torch.bool)
    for r in range(world):
        n = int(metas[r][0])
        rows = id_list[r][:n]
        if n and (bool((rows < 0).any()) or bool((rows >= n_total).any())):
            raise RuntimeError("gather_padded: rank %d sent an index outside [0, %d)" % (r, n_total))
        if bool(seen[rows.cpu()].any()):
            raise RuntimeError("gather_padded: an utterance was produced by two ranks")
        seen[rows.cpu()] = True
        out[rows] = buf_list[r][:n]
    if not bool(seen.all()):
        raise RuntimeError("gather_padded: some utterances were produced by no rank")
    return out


def synthesise_sharded(fn: Callable[..., torch.Tensor], lengths: Sequence[int], *batched: torch.Tensor, dst: int = 0,
                       group: Optional[dist.ProcessGroup] = None) -> Optional[torch.Tensor]:
    """Run `fn` on this rank's share of the batch and gather the result on `dst`.

    `batched` are tensors whose first dimension indexes utterances (mu, mask, noise, speaker ids ...); `fn` receives
    their rows for this rank (same order as `partition` assigns them) and returns (b_local, C, L).  With an empty share
    `fn` is not called.  Without an initialised process group this is plain single-process execution.
    """
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    idx = partition(lengths, world)[rank]
    if idx:
        sel = torch.as_tensor(idx, dtype=torch.long)
        local = fn(*[t[sel.to(t.device)] for t in batched])
    else:  # nothing to do on this rank: contribute an empty block of the right rank / dtype / device
        ref = batched[0]
        local = ref.new_zeros(0, 1, 1)
    if world > 1:
        # channel counts must agree for the padded gather; a rank with an empty share learns them from the others
        ch = torch.tensor([local.shape[1] if local.shape[0] else 0], dtype=torch.long, device=local.device)
        chs = [torch.empty_like(ch) for _ in range(world)]
        dist.all_gather(chs, ch, group=group)
        if local.shape[0] == 0:
            local = local.new_zeros(0, max(int(c) for c in chs), 1)
    return gather_padded(local, idx, len(lengths), dst=dst, group=group)


def allreduce_gradients(params, bucket_bytes: int = 32 << 20, group: Optional[dist.ProcessGroup] = None,
                        valid_frames: Optional[torch.Tensor] = None) -> None:
    """Data-parallel training (BASELINE cfg5): average the estimator's gradients over ranks after CFM.compute_loss's
    backward.  This is the training path's one real exchange step: gradients are flattened into
    buckets of at most `bucket_bytes` (11.0 M fp32 parameters = 44 MB -> two 32 MB buckets, sized for launch latency,
    not link count: NVSwitch reduces in the switch), each bucket is one all-reduce (NCCL on GPUs, gloo in the CPU
    tests) issued asynchronously so the copies of bucket i+1 overlap the reduction of bucket i.

    The reference trains through Lightning DDP (configs/trainer/ddp.yaml), whose default is the plain mean over ranks;
    pass `valid_frames` (this rank's sum(mask), a 0-dim tensor) to weight ranks by their valid frames instead, which
    makes the result equal the single-process gradient of the concatenated batch (the loss is normalised by
    sum(mask) * 80, flow_matching.py:115-117)."""
    params = [p for p in params if p.grad is not None]
    if not params or not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    world = dist.get_world_size(group)
    weight = None
    if valid_frames is not None:   # stays on the device: no host synchronisation between the backward and the exchange
        total = valid_frames.detach().clone().to(params[0].grad.device, torch.float32).reshape(1)
        mine = total.clone()
        dist.all_reduce(total, group=group)
        weight = mine / total
    buckets, cur, size = [], [], 0
    for p in params:
        n = p.grad.numel() * p.grad.element_size()
        if cur and size + n > bucket_bytes:
            buckets.append(cur)
            cur, size = [], 0
        cur.append(p)
        size += n
    if cur:
        buckets.append(cur)
    pending = []
    for bucket in buckets:
        flat = torch.cat([p.grad.reshape(-1) for p in bucket])
        if weight is not None:
            flat.mul_(weight)
        pending.append((bucket, flat, dist.all_reduce(flat, group=group, async_op=True)))
    for bucket, flat, work in pending:
        work.wait()
        if weight is None:
            flat.div_(world)
        # one batched copy back into the 180 gradient tensors instead of one small kernel each
        views = [v.view_as(p.grad) for v, p in zip(flat.split([p.grad.numel() for p in bucket]), bucket)]
        torch._foreach_copy_([p.grad for p in bucket], views)
