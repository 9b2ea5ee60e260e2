"""Data-parallel decoding: one process per GPU, whole reference batches sharded statically, ONE collective at
the end.  Replaces the reference's threaded DataParallelModel (reference src/parallel.py:70-148: per-call
parameter re-broadcast + scatter, outputs left as per-GPU Python lists, src/translate.py:50-54).

Sharding whole batches (never re-batching) keeps every sentence's result identical for any GPU count, because the
search of a batch runs until ALL its rows have finished (reference src/seq_gen.py:134).
"""
from typing import List, Sequence

import torch
import torch.distributed as dist


def batch_cost(n_sent: int, src_len: int, beam: int, max_len_a: float = 1.3, max_len_b: int = 5) -> float:
    """Relative cost model T * (c0 + c1*M + c2*M*T/2): weight streaming per step + per-row work + KV reads."""
    T = min(int(max_len_a * src_len + max_len_b), 512)
    M = n_sent * beam
    return T * (30.0 + 0.12 * M + 0.0028 * M * (T / 2.0 + src_len))


def shard_batches(costs: Sequence[float], world_size: int) -> List[List[int]]:
    """Greedy longest-processing-time assignment of batch indices to ranks (deterministic)."""
    order = sorted(range(len(costs)), key=lambda i: (-costs[i], i))
    load = [0.0] * world_size
    shards = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        shards[r].append(i)
        load[r] += costs[i]
    for s in shards:
        s.sort()
    return shards


def gather_outputs(ids: Sequence[int], outputs: Sequence[torch.Tensor], device, dst: int = 0):
    """Gathers (sentence id, tokens) pairs from all ranks onto rank `dst` with ONE padded all_gather
    (NCCL over NVLink on GPUs, gloo in the CPU tests).  Returns {id: 1-D LongTensor} on dst, None elsewhere."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    if world == 1:
        return {int(i): o for i, o in zip(ids, outputs)}
    n = len(outputs)
    width = max([int(o.numel()) for o in outputs] + [0])
    meta = torch.tensor([n, width], dtype=torch.int64, device=device)
    dist.all_reduce(meta, op=dist.ReduceOp.MAX)
    n_max, w_max = int(meta[0]), int(meta[1])
    # one row per sentence: [id, length, tokens...]; id = -1 marks padding rows
    pack = torch.full((n_max, w_max + 2), -1, dtype=torch.int32)
    for r, (i, o) in enumerate(zip(ids, outputs)):
        pack[r, 0], pack[r, 1] = int(i), int(o.numel())
        pack[r, 2:2 + o.numel()] = o.to(torch.int32)
    pack = pack.to(device)
    parts = [torch.empty_like(pack) for _ in range(world)]
    dist.all_gather(parts, pack)
    if dist.get_rank() != dst:
        return None
    res = {}
    allp = torch.cat(parts, 0).cpu()
    for row in allp:
        if int(row[0]) >= 0:
            res[int(row[0])] = row[2:2 + int(row[1])].to(torch.int64)
    return res
