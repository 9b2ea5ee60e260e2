"""Multi-GPU behaviour (-m gpu; every test skips on a box with one device -- run with `gpurun --gpus 2`, logs under profiles/).

  * the reference's own multi-GPU mode: ONE process, one Python thread per GPU (src/parallel.py:104-148 parallel_apply under
    DataParallelModel) -- nothing in the library may be cached per process where CUDA keeps it per device;
  * one process per GPU over NCCL (this package's mode): a corpus sharded by whole batches over 2 ranks and
    gathered must equal the single-rank decode of the same corpus, sentence for sentence (BASELINE.json configs[4]).
"""
import os
import socket
import threading

import pytest
import torch
import torch.multiprocessing as mp

from imagetranslate_b200 import BeamDecoder
from oracle import cases, restated as R

pytestmark = pytest.mark.gpu
two_gpus = pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two CUDA devices")


def _model_and_batches(seed=91, V=3000, n_batches=6):
    tp = R.FakeTextProcessor(V)
    torch.manual_seed(seed)
    model = R.Seq2SeqModel(tp, lang_dec=False, tie_embed=False, enc_layer=1, dec_layer=2).eval()
    with torch.no_grad():
        model.output_layer[1].layer.bias[4] += 0.8
    batches = []
    for i in range(n_batches):
        B, S = 3 + (i * 5) % 7, 8 + 3 * i
        src, mask, pad_idx = cases.make_text_batch(B, S, V, 700 + i)
        batches.append(dict(src_inputs=src, src_mask=mask, src_sizes=pad_idx, src_langs=torch.zeros(B, dtype=torch.long),
                            first_tokens=torch.full((B,), 6), tgt_langs=torch.ones(B, dtype=torch.long), pad_idx=0))
    return model, batches


@two_gpus
@pytest.mark.parametrize("dtype", ["fp32", "fp16"])
def test_one_process_two_devices_two_threads(dtype):
    """Decodes the same batches on cuda:0 and cuda:1 from two threads of one process (the first tcgen05 launch of every kernel
    happens on BOTH devices: MaxDynamicSharedMemorySize is a per-device function attribute) and gets identical tokens."""
    model, batches = _model_and_batches()
    import copy
    gens = []
    for dev in ("cuda:0", "cuda:1"):
        m = copy.deepcopy(model).to(dev)
        gens.append(BeamDecoder(m, beam_width=4, dtype=dtype, **cases.GEN))
    outs, errs = [None, None], []

    def work(i):
        try:
            with torch.cuda.device(i), torch.no_grad():
                res = [gens[i](**kw) for kw in batches]
                res.append(gens[i].forward_many(batches, lanes=2, rows=64))
                torch.cuda.synchronize()
            outs[i] = res
        except Exception as e:                                    # surfaced below, on the test's thread
            errs.append((i, repr(e)))

    th = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs
    flat = lambda r: [[o.tolist() for o in b] for b in r[:-1]] + [[o.tolist() for o in b] for b in r[-1]]
    assert flat(outs[0]) == flat(outs[1])
    if dtype == "fp32":
        want = [[w.tolist() for w in R.beam_search(model, beam_width=4, **cases.GEN, **kw)] for kw in batches[:2]]
        assert flat(outs[1])[:2] == want


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, dtype, q):
    import torch.distributed as dist
    from imagetranslate_b200.parallel import batch_cost, gather_outputs, shard_batches
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        model, batches = _model_and_batches(n_batches=9)
        gen = BeamDecoder(model.to("cuda:%d" % rank), beam_width=4, dtype=dtype, **cases.GEN)
        costs = [batch_cost(int(b["src_inputs"].size(0)), int(b["src_inputs"].size(1)), 4) for b in batches]
        mine = shard_batches(costs, world)[rank]
        first, o = [], 0
        for b in batches:
            first.append(o)
            o += int(b["src_inputs"].size(0))
        ids, outs = [], []
        with torch.no_grad():
            for bi, res in zip(mine, gen.forward_many([batches[i] for i in mine], lanes=2, rows=64)):
                ids += [first[bi] + k for k in range(len(res))]
                outs += res
        got = gather_outputs(ids, outs, torch.device("cuda", rank))
        if rank == 0:
            q.put([t.tolist() for t in got.in_order()])
    finally:
        dist.destroy_process_group()


@two_gpus
@pytest.mark.parametrize("dtype", ["fp32", "fp16"])
def test_two_rank_sharded_decode_equals_single_rank(dtype):
    model, batches = _model_and_batches(n_batches=9)
    gen = BeamDecoder(model.to("cuda:0"), beam_width=4, dtype=dtype, **cases.GEN)
    with torch.no_grad():
        single = [o.tolist() for b in gen.forward_many(batches, lanes=2, rows=64) for o in b]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, dtype, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert got == single


@two_gpus
def test_translate_driver_under_torchrun_writes_the_single_gpu_file(tmp_path):
    """python -m imagetranslate_b200.translate on 2 ranks (whole batches assigned by cost, ONE gather, rank 0 writes) must
    write the same file as the single-process run (BASELINE configs[0] tokenizer / input, a small seeded model)."""
    import subprocess
    import sys
    import types
    from imagetranslate_b200 import translate as tr
    from imagetranslate_b200.models import Seq2Seq
    from imagetranslate_b200.textprocessor import TextProcessor
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "config1")
    tp = TextProcessor(gold)
    torch.manual_seed(3)
    model = Seq2Seq(tp, lang_dec=False, tie_embed=False, enc_layer=1, dec_layer=1)
    with torch.no_grad():
        model.output_layer[0].layer.bias[4] += 1.0
    model_dir = str(tmp_path / "model")
    model.save(model_dir)
    common = ["--tok", gold, "--model", model_dir, "--input", os.path.join(gold, "input.txt"), "--src", "fa", "--target", "en",
              "--beam", "4", "--batch", "400", "--capacity", "600"]          # --batch 400: several batches to shard
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    one = str(tmp_path / "one.txt")
    r = subprocess.run([sys.executable, "-m", "imagetranslate_b200.translate", "--output", one] + common, cwd=root,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    two = str(tmp_path / "two.txt")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", str(_free_port()), "-m", "imagetranslate_b200.translate", "--output", two] + common,
                       cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    a, b = open(one).read(), open(two).read()
    assert len(a) > 100 and a == b
