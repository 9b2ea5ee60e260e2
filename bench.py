#!/usr/bin/env python
"""Benchmark of the beam-search decode path: decoded tokens/s at beam 4 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W]            this repo's CUDA path, headline config (configs[1])
  python bench.py --impl reference ...                           the reference algorithm on the host CPU cores
  python bench.py --config 3|4|5 ...                             the other BASELINE.json configs (captioning, S=256 beam sweep,
                                                                 length-bucketed corpus sharded over the ranks)

Headline workload = BASELINE.json configs[1]: MassSeq2Seq base, vocab 60k, 6+6 layers, d=768, synthetic sentences of length
16..128, length-sorted and batched by the reference's token-budget rule (--batch 4000 --capacity 600), beam 4.
One "step" = one full decode pass (encoder memory + every search step + output extraction [+ the NCCL gather at N > 1]) over
the step's sentence set.  Weak scaling: every rank decodes its own sentence set of the same size.

dtype: the model's weights are bf16 (model.bfloat16()); the kernels hold them exactly in fp16 and run fp16 x fp16 -> fp32 on
tcgen05 (kind::f16), with fp16 activations / KV cache and fp32 softmax, LayerNorm, logits and scores -- the arithmetic that
meets the north_star parity bar (tests/test_gpu_parity16.py).  --dtype bf16 runs pure bf16 arithmetic (same speed).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

VOCAB, ENC, DEC, BEAM = 60000, 6, 6, 4
GEN = dict(max_len_a=1.3, max_len_b=5, len_penalty_ratio=0.8)     # reference src/translate.py:26-28


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1, help="ranks the launcher started (informational; the world size comes from torchrun's env)")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=[2, 3, 4, 5], help="BASELINE.json configs (1-based); 2 = headline")
    ap.add_argument("--sentences", type=int, default=None, help="sentences per rank per step (config 2: 4096; config 5: 131072)")
    ap.add_argument("--batch", type=int, default=4000)
    ap.add_argument("--capacity", type=int, default=600)
    ap.add_argument("--dtype", default="fp16", choices=["fp16", "bf16", "fp32"])
    ap.add_argument("--rows", type=int, default=4096, help="beam rows per super-batch (consecutive reference batches decoded in lockstep)")
    ap.add_argument("--lanes", type=int, default=3, help="super-batches in flight per GPU (separate streams)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the per-batch / GPU-eager / extra roofline legs")
    ap.add_argument("--cpu-sentences", type=int, default=12)
    ap.add_argument("--eos-bias", type=float, default=0.0, help="added to the EOS logit bias: 1.5-2.5 makes a share of the hypotheses finish early (random-init weights never emit EOS on their own, so the default workload runs every sentence to its max_len; 6 ends every sentence at step 1)")
    ap.add_argument("--cpu-full", action="store_true", help="cpu baseline on the FULL median batch (minutes)")
    ap.add_argument("--single-leg", action="store_true", help="time ONE leg only, from pinned host inputs (value = e2e); for the long config-5 passes")
    ap.add_argument("--no-gather", action="store_true", help="diagnostic: leave the NCCL output gather out of the passes (N > 1)")
    return ap.parse_args()


def build_batches(n_sent, seed, batch, capacity, min_len=16, max_len=128, body_seed=None):
    from imagetranslate_b200.dataset import MTDataset, synthetic_sentences
    sents = synthetic_sentences(n_sent, VOCAB, min_len, max_len, seed=seed, body_seed=body_seed)
    ex = [(s, [6], 0, 1) for s in sents]
    return MTDataset(examples=ex, max_batch_capacity=capacity, max_batch=batch, pad_idx=0, max_seq_len=10000).batches


def batch_kwargs(b, device=None, pin=False):
    kw = dict(src_inputs=b["src_texts"], src_sizes=b["pad_idx"], first_tokens=b["dst_texts"][:, 0],
              src_mask=b["src_pad_mask"], src_langs=b["src_langs"], tgt_langs=b["dst_langs"])
    if device is not None:
        kw = {k: (v.to(device) if k != "src_sizes" else v) for k, v in kw.items()}
    elif pin:
        kw = {k: v.pin_memory() for k, v in kw.items()}
    kw["pad_idx"] = 0
    return kw


class ClockSampler(object):
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", os.environ.get("BENCH_CLOCK_MS", "200")], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        load = [x for x in sm if mx and x > 0.5 * mx] or sm
        med = load[len(load) // 2] if load else None
        return {"sm_mhz": med, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def emitted(outs):
    return sum(o.shape[0] for o in outs) - len(outs)   # language token excluded, as translate.py:55 decodes x[1:]


# ------------------------------------------------------------------------------------------- reference arms (oracle port)
def _oracle_model(dtype=torch.float32, device="cpu"):
    from oracle import restated as R
    tp = R.FakeTextProcessor(VOCAB)
    torch.manual_seed(0)
    model = R.Seq2SeqModel(tp, lang_dec=False, tie_embed=False, enc_layer=ENC, dec_layer=DEC).eval()
    return model.to(device=device, dtype=dtype)


def _median_batch(args, n=None):
    bs = build_batches(4096, 1234, args.batch, args.capacity)
    b = bs[len(bs) // 2]                 # the median-length batch of the workload (the first one holds the shortest sentences)
    if n is not None:
        b = {k: (v[:n] if v.dim() > 0 and v.size(0) >= n else v) for k, v in b.items()}
    return b


def run_reference(args, rank):
    """Reference arm / cpu_baseline: the reference algorithm (oracle port of seq_gen.py BeamDecoder, full-prefix
    recompute, fp32) on the host cores, on a bounded sample of the same workload."""
    from oracle import restated as R
    torch.set_num_threads(os.cpu_count() or 1)
    model = _oracle_model()
    full = _median_batch(args)
    n = int(full["src_texts"].size(0)) if args.cpu_full else min(args.cpu_sentences, int(full["src_texts"].size(0)))
    b = _median_batch(args, n)
    kw = batch_kwargs(b)
    sample = "%s %d sentences of the median-length reference batch of configs[1] (%d sentences, src len %d; full max_len), fp32, %d threads" % (
        "all" if args.cpu_full else "first", n, int(full["src_texts"].size(0)), b["src_texts"].size(1), torch.get_num_threads())
    times, toks = [], 0
    for it in range(args.warmup + args.steps):
        t0 = time.time()
        outs = R.beam_search(model, beam_width=BEAM, **GEN, **kw)
        dt = time.time() - t0
        if it >= args.warmup:
            times.append(dt)
            toks = emitted(outs)
    ms = 1000.0 * sum(times) / len(times)
    return dict(value=toks / (ms / 1000.0), unit="tokens/s", cores=torch.get_num_threads(), kind="port", sample=sample,
                ms_per_step=ms)


def gpu_eager_baseline(args, dev):
    """The reference's own single-GPU PyTorch BeamDecoder (north_star's ">= 20x" denominator): the oracle port of
    src/seq_gen.py -- plain PyTorch eager, full-prefix recompute -- on this GPU, fp32 and .bfloat16() / .half() (the stand-ins
    for apex O2), on the FULL median-length reference batch of configs[1]."""
    from oracle import restated as R
    b = _median_batch(args)
    res = {}
    for name, dt in (("fp32", torch.float32), ("bf16", torch.bfloat16), ("fp16", torch.float16)):
        try:
            model = _oracle_model(dt, dev)
            kw = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in batch_kwargs(b).items()}
            best = None
            for _ in range(2):
                torch.cuda.synchronize()
                t0 = time.time()
                with torch.no_grad():
                    outs = R.beam_search(model, beam_width=BEAM, **GEN, **kw)
                torch.cuda.synchronize()
                dtm = time.time() - t0
                best = dtm if best is None else min(best, dtm)
            res[name] = dict(value=emitted(outs) / best, unit="tokens/s", ms=1000.0 * best)
            del model
            torch.cuda.empty_cache()
        except Exception as e:                     # a dtype the eager path cannot run is reported, not hidden
            res[name] = dict(error=repr(e)[:200])
    res["sample"] = "full median-length reference batch of configs[1]: %d sentences, src len %d, beam 4" % (
        int(b["src_texts"].size(0)), int(b["src_texts"].size(1)))
    res["kind"] = "oracle port of the reference BeamDecoder (plain PyTorch eager, full-prefix recompute) on 1 B200"
    return res


# ------------------------------------------------------------------------------------------- kernel rooflines
def _time_launches(fn, n=20, flush=None):
    """Average duration of single launches, each timed on its own with CUDA events on the launching stream, L2 flushed
    (a 256 MB write) between them: the protocol of round 1's `roofline` entry, kept for the fused vocabulary projection so
    that the figure stays comparable across rounds (a kernel timed alone -> burst peak)."""
    for _ in range(3):
        fn()
    tot = 0.0
    for _ in range(n):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / n


def _time_rotating(fn, n_sets, reps=3):
    """Average duration of one launch of fn(i), i cycling over `n_sets` operand sets whose cold part (weights / K,V cache)
    adds up to more than the 126 MB L2, launched back to back on the current stream and timed with CUDA events around
    the whole train (>= 2 warm-up trains first).  Every launch therefore streams its weights / cache from HBM, as in a
    search step, while its small activation operands stay L2-resident, as in a search step."""
    for _ in range(2):
        for i in range(n_sets):
            fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps):
        for i in range(n_sets):
            fn(i)
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / (reps * n_sets)


def time_kernel_rooflines(gen, model, peaks, rows, code, extras=True):
    """Times the dominant kernels of the search step at the super-batch row count with CUDA events on the launching stream:
    the fused vocabulary projection as single launches with L2 flushed in between (round 1's protocol), the others as trains
    of back-to-back launches over rotating operand sets larger than L2 (see _time_rotating):
      roofline              the fused vocabulary projection (tcgen05 GEMM whose epilogue emits the softmax / top-k partials),
                            tensor-bound: algorithmic flops = 2 * rows * V * d
      roofline_gemm_layers  the four layer-GEMM shapes of a decoder layer (QKV, O = CQ = CO, FFN1 + GELU, FFN2), tensor
      roofline_attention    cross-attention over S = 100 memory positions and self-attention at t = 80, HBM-bound:
                            algorithmic bytes = K,V rows that must be read once"""
    import ctypes as C
    from imagetranslate_b200 import _lib
    dev = torch.device("cuda", torch.cuda.current_device())
    pw = gen._weights(model, 1, code, dev)
    lib = _lib.lib()
    act = torch.float16 if code == _lib.F16 else torch.bfloat16
    M, d, V, F, H = rows, pw.d, pw.V, pw.F, pw.H
    nc, KT = lib.itb_vocab_chunks(V), lib.itb_topk_capacity(BEAM)
    x = torch.randn(M, d, device=dev).to(act)
    cand_val = torch.empty(M, nc, KT, device=dev)
    cand_idx = torch.empty(M, nc, KT, dtype=torch.int32, device=dev)
    cmax, csum = torch.empty(M, nc, device=dev), torch.empty(M, nc, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    peak = peaks.get("bf16_tflops", 1590.0)
    hbm = peaks.get("hbm_gbs", 6550.0)
    copies = lambda t, total_mb: [t] + [t.clone() for _ in range(max(1, int(total_mb * 1e6 / (t.numel() * t.element_size()))) )]

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def vocab():
        _lib.check(lib.itb_vocab_partials(code, BEAM, _lib.ptr(x), _lib.ptr(pw.w_vocab), _lib.ptr(pw.b_vocab), M, V, d,
                                          _lib.ptr(cand_val), _lib.ptr(cand_idx), _lib.ptr(cmax), _lib.ptr(csum), None, stream))

    ms = _time_launches(vocab, 20, flush)
    del flush
    tf = 2.0 * M * V * d / ms / 1e9
    traffic = None
    try:        # dram__bytes_read.sum + dram__bytes_write.sum of this kernel at this shape, from the committed ncu capture
        tj = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))
        if tj.get("rows") == M and tj.get("V") == V:
            traffic = tj["dram_bytes_per_launch"]
    except Exception:
        pass
    alg = V * d * 2 + M * d * 2 + M * nc * (2 + 2 * KT) * 4
    out = {"roofline": dict(bound="tensor", achieved=tf, peak=peak, unit="TFLOP/s", frac=tf / peak, traffic=traffic,
                            kernel="gemm_tc_kernel<256, vocab-partials> rows=%d V=%d K=%d (fused logits + softmax/top-k partials)" % (M, V, d),
                            ms=ms, algorithmic_bytes=alg, hbm_gbs_at_algorithmic_bytes=alg / ms / 1e6,
                            protocol="single launches, CUDA events on the launching stream, L2 flushed between launches")}
    if not extras:
        return out
    # ---- layer GEMMs: weights rotate over > 160 MB of copies (cold, as between two search steps), activations stay hot
    gl = {}
    ly = pw.keep[0]
    y = torch.empty(_lib.Y_PARTIALS * M * max(3 * d, F), dtype=torch.float32, device=dev)
    hbuf = torch.randn(M, F, device=dev).to(act)
    for name, w, b, a, N, K, epi in (("qkv", ly["w_qkv"], ly["b_qkv"], x, 3 * d, d, _lib.EPI_BIAS_F32),
                                     ("o_cq_co", ly["w_o"], ly["b_o"], x, d, d, _lib.EPI_BIAS_F32),
                                     ("ffn1_gelu", ly["w_1"], ly["b_1"], x, F, d, _lib.EPI_BIAS_GELU_ACT),
                                     ("ffn2", ly["w_2"], ly["b_2"], hbuf, d, F, _lib.EPI_BIAS_F32)):
        ws_ = copies(w, 160)
        fn = lambda i, ws_=ws_, b=b, a=a, N=N, K=K, epi=epi: _lib.check(lib.itb_linear(code, _lib.GEMM_AUTO, epi, _lib.ptr(a), _lib.ptr(ws_[i]),
                                                                                       _lib.ptr(b), _lib.ptr(y), M, N, K, N, stream))
        t = _time_rotating(fn, len(ws_), 2)
        f = 2.0 * M * N * K / t / 1e9
        gl[name] = dict(ms=t, achieved=f, frac=f / peak, shape="M=%d N=%d K=%d" % (M, N, K), weight_copies=len(ws_))
        del ws_
    tot_ms = gl["qkv"]["ms"] + 3 * gl["o_cq_co"]["ms"] + gl["ffn1_gelu"]["ms"] + gl["ffn2"]["ms"]
    tot_fl = 2.0 * M * (6 * d * d + 2 * d * F)
    out["roofline_gemm_layers"] = dict(bound="tensor", unit="TFLOP/s", peak=peak, achieved=tot_fl / tot_ms / 1e9,
                                       frac=tot_fl / tot_ms / 1e9 / peak, per_shape=gl,
                                       note="one decoder layer's six GEMMs (QKV + O + CQ + CO + FFN1 + FFN2); each shape a train of launches over rotating weight copies > L2")
    # ---- attention (HBM): 4 rotating K,V buffers of 0.3 GB each
    B, S, t_pos, T = M // BEAM, 100, 80, 96
    esz = 2
    ctl = torch.tensor([t_pos, 0, t_pos - 1, 0, 0, 0, 1000, 0], dtype=torch.int32, device=dev)
    qkv = torch.randn(M, 3 * d, device=dev)
    ckv = [torch.randn(B, S, 1, 2, d, device=dev).to(act) for _ in range(4)]
    mask = torch.ones(B, S, dtype=torch.uint8, device=dev)
    ctx = torch.empty(M, d, dtype=act, device=dev)
    cross = lambda i: _lib.check(lib.itb_cross_attention(code, _lib.ptr(ctl), None, None, _lib.ptr(qkv), 3 * d, _lib.ptr(ckv[i]), 0, _lib.ptr(mask),
                                                         _lib.ptr(ctx), M, BEAM, d, H, S, 1, stream))
    t_c = _time_rotating(cross, 4, 4)
    by_c = B * S * 2 * d * esz
    del ckv
    kv = [torch.randn(M, T, 2, d, device=dev).to(act) for _ in range(2)]         # 1.2 GB each
    anc = torch.randint(0, BEAM, (2, M, T), device=dev)
    share = t_pos - 6                  # hypotheses of a sentence share their prefix up to the last few positions
    anc[:, :, :share] = anc.view(2, B, BEAM, T)[:, :, :1, :share].expand(2, B, BEAM, share).reshape(2, M, share)
    anc = anc.to(torch.uint8)
    self_ = lambda i: _lib.check(lib.itb_self_attention(code, _lib.ptr(ctl), None, _lib.ptr(qkv), _lib.ptr(kv[i]), _lib.ptr(anc), _lib.ptr(ctx),
                                                        M, BEAM, d, H, T, stream))
    t_s = _time_rotating(self_, 2, 6)
    by_s = (B * share + M * (t_pos - share)) * 2 * d * esz
    out["roofline_attention"] = dict(bound="hbm", unit="GB/s", peak=hbm, achieved=by_c / t_c / 1e6, frac=by_c / t_c / 1e6 / hbm,
                                     kernel="cross-attention (attn_bulk_kernel: cp.async.bulk + mbarrier ring), %d sentences x S=%d, one layer" % (B, S),
                                     ms=t_c, algorithmic_bytes=by_c, protocol="train of launches over 4 rotating K,V buffers (1.26 GB > L2)",
                                     self_attention=dict(ms=t_s, algorithmic_bytes=by_s, achieved=by_s / t_s / 1e6, frac=by_s / t_s / 1e6 / hbm,
                                                         kernel="self-attention (attn_mma_kernel) t=%d, %d rows, prefix shared up to position %d" % (t_pos, M, share)))
    return out


# ------------------------------------------------------------------------------------------- the decode arms
def _dist_max_sum(vals, dev, world):
    import torch.distributed as dist
    t = torch.tensor(vals, dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        return tmax.tolist(), tsum.tolist()
    return t.tolist(), t.tolist()


def _timed(one_pass, steps, warmup, dev, world, sampler=None):
    """W warm-up passes, then K passes bracketed by barrier + synchronize, CUDA events, max over ranks."""
    import gc
    import torch.distributed as dist
    for _ in range(warmup):
        one_pass()
    gc.collect()           # start the timed region with a clean heap
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    toks = launches = 0
    for _ in range(steps):
        t, nl = one_pass()
        toks += t
        launches += nl
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    mx, sm = _dist_max_sum([ms, float(toks), float(launches)], dev, world)
    return mx[0], sm[1], sm[2], clocks


def make_text_model(args, dev):
    from imagetranslate_b200 import BeamDecoder
    from imagetranslate_b200.models import MassSeq2Seq, SyntheticTextProcessor
    torch.manual_seed(0)
    model = MassSeq2Seq(SyntheticTextProcessor(VOCAB), lang_dec=False, tie_embed=False, enc_layer=ENC, dec_layer=DEC).eval()
    if args.eos_bias != 0.0:
        with torch.no_grad():
            for ol in model.output_layer:
                ol.layer.bias[4] += args.eos_bias
    model = model.to(dev)
    if args.dtype != "fp32":
        model = model.to(torch.bfloat16)
    return model, BeamDecoder(model, beam_width=BEAM, dtype=args.dtype, **GEN)


def run_text(args, rank, world, local_rank):
    """configs[1] (headline) and configs[4] (corpus sharded by cost over the ranks)."""
    from imagetranslate_b200 import _lib
    from imagetranslate_b200.parallel import batch_cost, gather_outputs, shard_batches
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    model, gen = make_text_model(args, dev)
    code = {"fp16": _lib.F16, "bf16": _lib.BF16, "fp32": _lib.F32}[args.dtype]
    if args.config == 5:
        # ONE global length-sorted corpus (seed 99), batched once, whole batches assigned to ranks by cost (LPT)
        n_total = (args.sentences or 131072) * world
        batches = build_batches(n_total, 99, args.batch, args.capacity)
        costs = [batch_cost(int(b["src_texts"].size(0)), int(b["src_texts"].size(1)), BEAM) for b in batches]
        mine = shard_batches(costs, world)[rank]
        first_id, o = [], 0
        for b in batches:
            first_id.append(o)
            o += int(b["src_texts"].size(0))
        ids = [first_id[i] + k for i in mine for k in range(int(batches[i]["src_texts"].size(0)))]
        batches = [batches[i] for i in mine]
        workload = "configs[4]: %d synthetic sentences (seed 99) len 16-128, ONE global length-sorted list batched by --batch %d --capacity %d, whole batches assigned to %d rank(s) by cost (LPT); %d batches on rank 0" % (
            n_total, args.batch, args.capacity, world, len(batches))
    else:
        n_sent = args.sentences or 4096
        # every rank decodes its OWN sentences (token bodies drawn per rank) of the SAME length multiset: equal work per rank
        batches = build_batches(n_sent, 1234, args.batch, args.capacity, body_seed=None if rank == 0 else 5000 + rank)
        ids = [rank * n_sent + k for k in range(n_sent)]
        n_total = n_sent * world
        workload = "configs[1]: MassSeq2Seq base V=60000 6+6 d=768 beam 4, %d synthetic sentences/rank len 16-128 length-sorted (same length multiset, different sentences on every rank), --batch %d --capacity %d (%d batches/rank), random-init weights" % (
            n_sent, args.batch, args.capacity, len(batches))
    host_kw = [batch_kwargs(b, pin=True) for b in batches]
    dev_kw = [] if args.single_leg else [batch_kwargs(b, device=dev) for b in batches]
    h2d = sum(sum(v.numel() * v.element_size() for v in kw.values() if torch.is_tensor(v)) for kw in host_kw)
    gathered = {}

    def one_pass(kws):
        outs, n0 = [], gen.kernel_launches
        with torch.no_grad():
            for o in gen.forward_many(kws, lanes=args.lanes, rows=args.rows):
                outs += o
        n_launch = gen.kernel_launches - n0
        toks = emitted(outs)
        if world > 1 and not args.no_gather:           # the NCCL output gather is part of every pass at N > 1
            res = gather_outputs(ids, outs, dev)
            if rank == 0:
                gathered["n"] = len(res)
                gathered["tokens"] = res.emitted_tokens()
        return toks, n_launch

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if args.single_leg:
        ms, toks, launches, clocks = _timed(lambda: one_pass(host_kw), args.steps, args.warmup, dev, world, sampler)
        ms_e, toks_e = ms, toks
    else:
        ms, toks, launches, clocks = _timed(lambda: one_pass(dev_kw), args.steps, args.warmup, dev, world, sampler)
        ms_e, toks_e, _, _ = _timed(lambda: one_pass(host_kw), args.steps, 1, dev, world)
    if world > 1 and rank == 0 and not args.no_gather:
        assert gathered["n"] == n_total, "gather returned %d of %d sentences" % (gathered["n"], n_total)
        assert gathered["tokens"] * args.steps == int(toks_e), "gathered token count differs from the ranks' own"
    d2h = int(toks_e / max(args.steps, 1) / world) * 4
    line = {
        "metric": "decoded tokens/sec at beam 4", "value": toks / (ms / 1000.0), "unit": "tokens/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": workload, "beam": BEAM, "lanes": args.lanes, "superbatch_rows": args.rows,
                   "arithmetic": {"fp16": "bf16-valued model weights held exactly in fp16; fp16 x fp16 -> fp32 tcgen05 MMAs, fp16 activations / KV cache, fp32 softmax / LayerNorm / logits / scores",
                                  "bf16": "bf16 weights, activations and KV cache; fp32 accumulation, softmax, LayerNorm, logits, scores",
                                  "fp32": "fp32 parity kernels (CUDA cores)"}[args.dtype],
                   "eos_bias": args.eos_bias, "single_leg": bool(args.single_leg),
                   "value_includes": "encoder + cross-KV + every search step + extraction" + (" + NCCL output gather" if world > 1 else ""),
                   "l2": "working set > L2: 191 MB of 16-bit weights streamed every search step plus the KV cache",
                   "parallelism": "dp%d (whole batches per rank, one all-gather of outputs at the end of every pass)" % world},
        "e2e": {"value": toks_e / (ms_e / 1000.0), "unit": "tokens/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e / args.steps},
        "gpu_launches": int(launches), "clocks": clocks,
    }
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        extras = world == 1 and not args.no_extras
        if args.dtype != "fp32":
            line.update(time_kernel_rooflines(gen, model, peaks, args.rows, code, extras))
            line["roofline"]["peak_source"] = "MEASURED_PEAKS.json bf16_tflops (burst, kernel timed alone) / hbm_gbs" if peaks else "fallback 1590 TFLOP/s, 6550 GB/s"
        if extras and args.config == 2:
            # the reference driver's call pattern: generator(...) once per reference batch (src/translate.py:46-49)
            def per_batch():
                n0, t = gen.kernel_launches, 0
                with torch.no_grad():
                    for kw in host_kw:
                        t += emitted(gen(**kw))
                return t, gen.kernel_launches - n0
            ms_b, toks_b, _, _ = _timed(per_batch, 1, 1, dev, 1)
            line["forward_per_batch"] = {"value": toks_b / (ms_b / 1000.0), "unit": "tokens/s", "ms_per_step": ms_b,
                                         "note": "the same pass through BeamDecoder.forward, one call per reference batch (host buffers in, CPU lists out)"}
            line["gpu_eager_baseline"] = gpu_eager_baseline(args, dev)
        if world == 1 and not args.no_cpu_baseline:
            cb = run_reference(argparse.Namespace(**dict(vars(args), steps=1, warmup=0)), 0)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line))


def run_caption(args, rank, world, local_rank):
    """configs[2]: ImageCaptioning (ResNet-50 grid 7x7x2048 -> text decoder), beam 4, 256 synthetic 224x224 images per rank,
    max_len = 256 (src/caption.py:38-39): as batches of 16 (caption.py's default) through forward_many, and as one batch of 256."""
    from imagetranslate_b200 import BeamDecoder
    from imagetranslate_b200.models import ImageCaptioning, SyntheticTextProcessor
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    torch.manual_seed(0)
    model = ImageCaptioning(SyntheticTextProcessor(VOCAB), resnet_depth=3, lang_dec=False, tie_embed=False, enc_layer=ENC,
                            dec_layer=DEC, use_obj=False).eval().to(dev)
    if args.dtype != "fp32":
        model = model.half()
    gen = BeamDecoder(model, beam_width=BEAM, dtype=args.dtype, **GEN)
    n_img = args.sentences or 256
    g = torch.Generator().manual_seed(7 + rank)
    images = torch.randn(n_img, 3, 224, 224, generator=g)
    # random-init BatchNorm statistics give grid features of huge magnitude; scale the projection so the
    # memory has O(1) entries like a trained model's
    with torch.no_grad():
        probe = model.encode(images=images[:16].to(dev))[0].float()
        model.image_model.fc.weight.mul_(1.0 / max(float(probe.std()), 1e-6))
    mk = lambda lo, hi: dict(first_tokens=torch.full((hi - lo,), 6), images=images[lo:hi].pin_memory(),
                             tgt_langs=torch.ones(hi - lo, dtype=torch.long), pad_idx=0, max_len=256)
    small = [mk(i, min(i + 16, n_img)) for i in range(0, n_img, 16)]
    big = [mk(0, n_img)]

    def one_pass(kws):
        n0, outs = gen.kernel_launches, []
        with torch.no_grad():
            for o in gen.forward_many(kws, lanes=args.lanes, rows=args.rows):
                outs += o
        return emitted(outs), gen.kernel_launches - n0

    def front_end():
        with torch.no_grad():
            model.encode(images=images.to(dev, non_blocking=True))
        return 0, 0

    sampler = ClockSampler(local_rank) if rank == 0 else None
    ms, toks, launches, clocks = _timed(lambda: one_pass(small), args.steps, args.warmup, dev, world, sampler)
    ms_big, toks_big, _, _ = _timed(lambda: one_pass(big), args.steps, 1, dev, world)
    ms_fe, _, _, _ = _timed(front_end, args.steps, 1, dev, world)
    h2d = images.numel() * 4
    line = {"metric": "decoded tokens/sec at beam 4", "value": toks / (ms / 1000.0), "unit": "tokens/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": "configs[2]: ImageCaptioning ResNet-50 grid (7x7x2048) -> 6-layer text decoder V=60000, beam 4, %d synthetic 224x224 images/rank, max_len 256, batches of 16 decoded in lockstep (forward_many); images start in pinned host memory" % n_img,
                       "one_batch_of_%d" % n_img: {"tokens_per_s": toks_big / (ms_big / 1000.0), "ms_per_step": ms_big / args.steps},
                       "resnet_front_end_ms_per_%d_images" % n_img: ms_fe / args.steps, "beam": BEAM},
            "e2e": {"value": toks / (ms / 1000.0), "unit": "tokens/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(toks / args.steps / world) * 4,
                    "ms_per_step": ms / args.steps},
            "gpu_launches": int(launches), "clocks": clocks}
    if rank == 0:
        print(json.dumps(line))


def run_beam_sweep(args, rank, world, local_rank):
    """configs[3]: all sources 256 tokens long (B = 15 at --batch 4000, and B = 64), beam 1 / 4 / 8 / 12, T = 337 steps."""
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    model, gen = make_text_model(args, dev)
    from imagetranslate_b200.dataset import synthetic_sentences
    res = {}
    for B in (15, 64):
        sents = synthetic_sentences(B, VOCAB, 256, 256, seed=4321)
        src = torch.stack(sents)
        kw = dict(src_inputs=src.pin_memory(), src_sizes=torch.full((B,), 255), first_tokens=torch.full((B,), 6),
                  src_mask=torch.ones(B, 256, dtype=torch.bool).pin_memory(), src_langs=torch.zeros(B, dtype=torch.long),
                  tgt_langs=torch.ones(B, dtype=torch.long), pad_idx=0)
        for beam in (1, 4, 8, 12):
            def one_pass():
                n0 = gen.kernel_launches
                with torch.no_grad():
                    outs = gen(beam_width=beam, **kw)
                return emitted(outs), gen.kernel_launches - n0
            ms, toks, launches, _ = _timed(one_pass, max(args.steps, 1), 1, dev, 1)
            steps = list(gen._engines.values())[0].last_steps
            res["B%d_beam%d" % (B, beam)] = dict(rows=B * beam, tokens_per_s=toks / (ms / 1000.0), ms_per_pass=ms / max(args.steps, 1),
                                                 search_steps=steps, us_per_search_step=1000.0 * ms / max(args.steps, 1) / max(steps, 1))
    b4 = res["B15_beam4"]
    line = {"metric": "decoded tokens/sec at beam 4", "value": b4["tokens_per_s"], "unit": "tokens/s", "n_gpus": 1, "steps": args.steps,
            "warmup": 1, "ms_per_step": b4["ms_per_pass"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": "configs[3]: sources of 256 tokens, one reference batch per call through BeamDecoder.forward (B = 15 at --batch 4000; B = 64), beam 1/4/8/12, 337 steps",
                       "sweep": res}}
    if rank == 0:
        print(json.dumps(line))


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if rank != 0:
            return
        cb = run_reference(args, 0)
        print(json.dumps({
            "impl": "reference", "metric": "decoded tokens/sec at beam 4", "value": cb["value"], "unit": "tokens/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[1] bounded sample: " + cb["sample"]},
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        if args.config == 3:
            run_caption(args, rank, world, local_rank)
        elif args.config == 4:
            run_beam_sweep(args, rank, world, local_rank)
        else:
            run_text(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
