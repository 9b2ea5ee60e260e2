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
    padded token matrix (rows are views; nothing is unpacked per sentence until it is asked for)."""

    def __init__(self, ids, lengths, tokens):
        self.ids, self.lengths, self.tokens = ids, lengths, tokens          # (n,) int64, (n,) int64, (n, W) int64
        self._row = dict(zip(ids.tolist(), range(ids.numel())))
        self._len = lengths.tolist()

    def __len__(self):
        return len(self._row)

    def __contains__(self, i):
        return int(i) in self._row

    def __getitem__(self, i):
        r = self._row[int(i)]
        return self.tokens[r, :self._len[r]]

    def keys(self):
        return self._row.keys()

    def items(self):
        return ((i, self.tokens[r, :self._len[r]]) for i, r in self._row.items())

    def in_order(self):
        """Outputs sorted by sentence id (what a writer emits: the reference writes in input order, translate.py:114-120)."""
        return [self[i] for i in sorted(self._row)]

    def emitted_tokens(self):
        """sum(len(output) - 1): the language token is not an emitted token (translate.py:55 decodes x[1:])."""
        return int(self.lengths.sum()) - int(self.lengths.numel())


def gather_outputs(ids: Sequence[int], outputs: Sequence[torch.Tensor], device, dst: int = 0):
    """Gathers (sentence id, tokens) pairs from all ranks onto rank `dst` with ONE padded all-gather
    (NCCL over NVLink on GPUs, gloo in the CPU tests).  Returns a GatheredOutputs mapping on dst, None elsewhere.
    Packing and unpacking are vectorised: no per-sentence Python on the send side, none on the receive side either
    until a row is read (the row loops of round 1 cost more than the collective itself)."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    n = len(outputs)
    lens = torch.tensor([int(o.numel()) for o in outputs], dtype=torch.int64)
    width = int(lens.max()) if n else 0
    ids_t = torch.as_tensor(list(ids), dtype=torch.int64)
    if n:
        padded = torch.nn.utils.rnn.pad_sequence(list(outputs), batch_first=True, padding_value=-1).to(torch.int64)
    else:
        padded = torch.zeros((0, 0), dtype=torch.int64)
    if world == 1:
        return GatheredOutputs(ids_t, lens, padded)
    meta = torch.tensor([n, width], dtype=torch.int64, device=device)
    dist.all_reduce(meta, op=dist.ReduceOp.MAX)
    n_max, w_max = int(meta[0]), int(meta[1])
    # one row per sentence: [id, length, tokens...]; id = -1 marks padding rows.  Sentence ids travel as int32.
    pack = torch.full((n_max, w_max + 2), -1, dtype=torch.int32)
    if n:
        pack[:n, 0], pack[:n, 1] = ids_t.to(torch.int32), lens.to(torch.int32)
        pack[:n, 2:2 + width] = padded.to(torch.int32)
    if torch.device(device).type == "cuda":
        pack = pack.pin_memory().to(device, non_blocking=True)
    allp = torch.empty((world * n_max, w_max + 2), dtype=torch.int32, device=device)
    dist.all_gather_into_tensor(allp, pack)
    if dist.get_rank() != dst:
        return None
    allp = allp.cpu()
    allp = allp[allp[:, 0] >= 0]
    return GatheredOutputs(allp[:, 0].to(torch.int64), allp[:, 1].to(torch.int64), allp[:, 2:].to(torch.int64))
