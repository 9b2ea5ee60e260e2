"""N>1 host path on CPU: whole reference batches are sharded over ranks and ONE all_gather brings the outputs
back in sentence order (replaces reference src/parallel.py:70-148 + src/translate.py:50-54).  world_size 2, gloo."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from imagetranslate_b200.dataset import MTDataset, synthetic_sentences
from imagetranslate_b200.parallel import batch_cost, gather_outputs, shard_batches


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_decode(sent_id):
    """Deterministic stand-in for a decoded hypothesis (ragged lengths, incl. length 1 = only the language token)."""
    g = torch.Generator().manual_seed(1000 + sent_id)
    n = 1 + int(torch.randint(0, 9, (1,), generator=g))
    return torch.randint(5, 60000, (n,), generator=g)


def _worker(rank, world, port, n_batches_sizes, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        costs = [batch_cost(n, s, 4) for n, s in n_batches_sizes]
        shards = shard_batches(costs, world)
        first_id, o = [], 0
        for n, _ in n_batches_sizes:
            first_id.append(o)
            o += n
        ids, outs = [], []
        for bi in shards[rank]:
            for k in range(n_batches_sizes[bi][0]):
                ids.append(first_id[bi] + k)
                outs.append(_fake_decode(first_id[bi] + k))
        res = gather_outputs(ids, outs, torch.device("cpu"))
        if rank == 0:
            q.put({k: v.tolist() for k, v in res.items()})
        else:
            assert res is None
    finally:
        dist.destroy_process_group()


def test_shard_batches_is_a_partition_and_balanced():
    costs = [5.0, 1.0, 9.0, 3.0, 3.0, 7.0, 2.0]
    for world in (1, 2, 3, 8):
        shards = shard_batches(costs, world)
        assert sorted(i for s in shards for i in s) == list(range(len(costs)))
        loads = [sum(costs[i] for i in s) for s in shards]
        assert max(loads) - min(l for l in loads if l > 0 or world <= len(costs)) <= max(costs)
    assert shard_batches(costs, 2) == shard_batches(costs, 2)          # deterministic


def test_two_rank_gather_restores_sentence_order():
    sents = synthetic_sentences(300, 60000, 16, 128, seed=5)
    batches = MTDataset(examples=[(s, [6], 0, 1) for s in sents], max_batch_capacity=600, max_batch=4000, pad_idx=0,
                        max_seq_len=10000).batches
    sizes = [(int(b["src_texts"].size(0)), int(b["src_texts"].size(1))) for b in batches]
    assert sum(n for n, _ in sizes) == 300 and len(sizes) > 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, sizes, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(got) == list(range(300))
    for i in range(300):
        assert got[i] == _fake_decode(i).tolist()


def test_single_process_gather_is_identity():
    outs = [_fake_decode(i) for i in range(5)]
    res = gather_outputs(list(range(5)), outs, torch.device("cpu"))
    assert [res[i].tolist() for i in range(5)] == [o.tolist() for o in outs]
