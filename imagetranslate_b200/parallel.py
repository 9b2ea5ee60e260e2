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


class GatheredOutputs(object):
    """Result of gather_outputs on the destination rank: a read-only mapping {sentence id: 1-D LongTensor} over ONE
    padded int32 matrix of rows [id, length, tokens...] exactly as it came off the collective.  Nothing is copied,
    converted or indexed per sentence until it is asked for (the id index is built on first lookup)."""

    def __init__(self, rows):
        self.rows = rows                                  # (n, 2 + W) int32; id = -1 marks padding rows of short shards
        self._index = None

    def _idx(self):
        if self._index is None:
            ids = self.rows[:, 0]
            valid = torch.nonzero(ids >= 0).reshape(-1)
            self._index = dict(zip(ids[valid].tolist(), valid.tolist()))
        return self._index

    @property
    def ids(self):
        return torch.tensor(sorted(self._idx()), dtype=torch.int64)

    def __len__(self):
        return int((self.rows[:, 0] >= 0).sum()) if self.rows.numel() else 0

    def __contains__(self, i):
        return int(i) in self._idx()

    def __getitem__(self, i):
        r = self._idx()[int(i)]
        return self.rows[r, 2:2 + int(self.rows[r, 1])].to(torch.int64)

    def keys(self):
        return self._idx().keys()

    def items(self):
        return ((i, self[i]) for i in self._idx())

    def in_order(self):
        """Outputs sorted by sentence id (what a writer emits: the reference writes in input order, translate.py:114-120)."""
        return [self[i] for i in sorted(self._idx())]

    def emitted_tokens(self):
        """sum(len(output) - 1): the language token is not an emitted token (translate.py:55 decodes x[1:])."""
        if not self.rows.numel():
            return 0
        valid = self.rows[:, 0] >= 0
        return int(self.rows[:, 1][valid].sum()) - int(valid.sum())


def gather_outputs(ids: Sequence[int], outputs: Sequence[torch.Tensor], device, dst: int = 0):
    """Gathers (sentence id, tokens) pairs from all ranks onto rank `dst` with ONE padded all-gather
    (NCCL over NVLink on GPUs, gloo in the CPU tests).  Returns a GatheredOutputs mapping on dst, None elsewhere.
    Packing is vectorised and the receive side is lazy: no per-sentence Python on either side of the collective
    (the row loops of round 1 cost more than the collective itself)."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    n = len(outputs)
    if n:
        padded = torch.nn.utils.rnn.pad_sequence(list(outputs), batch_first=True, padding_value=0)
        lens = torch.tensor([o.shape[0] for o in outputs], dtype=torch.int32)
        width = int(padded.shape[1])
    else:
        padded, lens, width = torch.zeros((0, 0), dtype=torch.int64), torch.zeros((0,), dtype=torch.int32), 0
    n_max, w_max = n, width
    if world > 1:
        meta = torch.tensor([n, width], dtype=torch.int64, device=device)
        dist.all_reduce(meta, op=dist.ReduceOp.MAX)
        n_max, w_max = int(meta[0]), int(meta[1])
    # one row per sentence: [id, length, tokens...]; id = -1 marks padding rows.  Sentence ids travel as int32.
    pack = torch.empty((n_max, w_max + 2), dtype=torch.int32, pin_memory=(world > 1 and torch.device(device).type == "cuda"))
    pack[:n, 0] = torch.as_tensor(list(ids), dtype=torch.int32)
    pack[:n, 1] = lens
    pack[:n, 2:2 + width] = padded
    pack[:n, 2 + width:] = 0
    pack[n:, 0] = -1
    pack[n:, 1:] = 0
    if world == 1:
        return GatheredOutputs(pack)
    allp = torch.empty((world * n_max, w_max + 2), dtype=torch.int32, device=device)
    dist.all_gather_into_tensor(allp, pack.to(device, non_blocking=True))
    if dist.get_rank() != dst:
        return None
    return GatheredOutputs(allp.cpu())
