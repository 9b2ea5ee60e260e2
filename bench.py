#!/usr/bin/env python
"""Headline benchmark: decoded tokens/s at beam 4 (BASELINE.json metric) on configs[1]:
MassSeq2Seq base, vocab 60k, 6+6 layers, d=768, bf16, synthetic sentences of length 16..128, length-sorted and
batched by the reference's token-budget rule (--batch 4000 --capacity 600).

  python bench.py [--gpus N] [--steps K] [--warmup W]            this repo's CUDA path
  python bench.py --impl reference ...                           the reference algorithm on the host CPU cores
One "step" = one full decode pass (encoder memory + every search step + output extraction) over the step's
sentence set.  Weak scaling: every rank decodes its own sentence set of the same size.
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
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--sentences", type=int, default=4096, help="sentences per rank per step")
    ap.add_argument("--batch", type=int, default=4000)
    ap.add_argument("--capacity", type=int, default=600)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp16", "fp32"])
    ap.add_argument("--rows", type=int, default=4096, help="beam rows per super-batch (consecutive reference batches decoded in lockstep)")
    ap.add_argument("--lanes", type=int, default=2, help="batches in flight per GPU (separate streams; results do not depend on it)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sentences", type=int, default=4)
    return ap.parse_args()


def build_batches(n_sent, seed, batch, capacity):
    from imagetranslate_b200.dataset import MTDataset, synthetic_sentences
    sents = synthetic_sentences(n_sent, VOCAB, 16, 128, seed=seed)
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
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
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
    return sum(int(o.numel()) - 1 for o in outs)       # language token excluded, as translate.py:55 decodes x[1:]


def run_reference(args, rank):
    """Reference arm / cpu_baseline: the reference algorithm (oracle port of seq_gen.py BeamDecoder, full-prefix
    recompute, fp32) on the host cores, on a bounded sample of the same workload."""
    from oracle import restated as R
    torch.set_num_threads(os.cpu_count() or 1)
    tp = R.FakeTextProcessor(VOCAB)
    torch.manual_seed(0)
    model = R.Seq2SeqModel(tp, lang_dec=False, tie_embed=False, enc_layer=ENC, dec_layer=DEC).eval()
    bs = build_batches(args.sentences, 1234, args.batch, args.capacity)
    b = bs[len(bs) // 2]                 # the median-length batch of the workload (the first one holds the shortest sentences)
    n = args.cpu_sentences
    kw = batch_kwargs({k: (v[:n] if v.dim() > 0 and v.size(0) >= n else v) for k, v in b.items()})
    sample = "first %d sentences of the median-length batch (src len %d, full max_len), fp32, %d threads" % (
        n, b["src_texts"].size(1), torch.get_num_threads())
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


def time_kernel_rooflines(gen, model, peaks, rows):
    """Times the dominant kernel of the step -- the fused vocabulary projection (tcgen05 GEMM whose epilogue emits the
    softmax / top-k partials) -- ALONE at the super-batch row count, with CUDA events on the launching stream and L2
    flushed between launches.  At these row counts it is tensor-bound: algorithmic flops = 2 * rows * V * d."""
    import ctypes as C
    from imagetranslate_b200 import _lib
    dev = torch.device("cuda", torch.cuda.current_device())
    pw = gen._weights(model, 1, _lib.BF16, dev)
    lib = _lib.lib()
    M, d, V = rows, pw.d, pw.V
    nc, KT = lib.itb_vocab_chunks(V), lib.itb_topk_capacity(BEAM)
    x = torch.randn(M, d, device=dev).to(torch.bfloat16)
    cand_val = torch.empty(M, nc, KT, device=dev)
    cand_idx = torch.empty(M, nc, KT, dtype=torch.int32, device=dev)
    cmax, csum = torch.empty(M, nc, device=dev), torch.empty(M, nc, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def fn():
        _lib.check(lib.itb_vocab_partials(_lib.BF16, BEAM, _lib.ptr(x), _lib.ptr(pw.w_vocab), _lib.ptr(pw.b_vocab), M, V, d,
                                          _lib.ptr(cand_val), _lib.ptr(cand_idx), _lib.ptr(cmax), _lib.ptr(csum), None, stream))

    for _ in range(3):
        fn()
    tot, n = 0.0, 20
    for _ in range(n):
        flush.zero_()                       # L2 flush between timed launches
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    ms = tot / n
    tf = 2.0 * M * V * d / ms / 1e9
    peak = peaks.get("bf16_tflops", 1590.0)
    traffic = None
    try:        # dram__bytes_read.sum + dram__bytes_write.sum of this kernel at this shape, from the committed ncu capture
        tj = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))
        if tj.get("rows") == M and tj.get("V") == V:
            traffic = tj["dram_bytes_per_launch"]
    except Exception:
        pass
    return dict(bound="tensor", achieved=tf, peak=peak, unit="TFLOP/s", frac=tf / peak, traffic=traffic,
                kernel="gemm_tc_kernel<256, vocab-partials> rows=%d V=%d K=%d (fused logits + softmax/top-k partials)" % (M, V, d),
                ms=ms, algorithmic_bytes=V * d * 2 + M * d * 2 + M * nc * (2 + 2 * KT) * 4,
                hbm_gbs_at_algorithmic_bytes=(V * d * 2 + M * d * 2 + M * nc * (2 + 2 * KT) * 4) / ms / 1e6)


def run_b200(args, rank, world, local_rank):
    import torch.distributed as dist
    from imagetranslate_b200 import BeamDecoder
    from imagetranslate_b200.models import MassSeq2Seq, SyntheticTextProcessor
    from imagetranslate_b200.parallel import gather_outputs
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    torch.manual_seed(0)
    model = MassSeq2Seq(SyntheticTextProcessor(VOCAB), lang_dec=False, tie_embed=False, enc_layer=ENC, dec_layer=DEC).eval()
    model = model.to(dev)
    if args.dtype != "fp32":
        model = model.to(torch.bfloat16)
    gen = BeamDecoder(model, beam_width=BEAM, dtype=args.dtype, **GEN)
    batches = build_batches(args.sentences, 1234 + rank, args.batch, args.capacity)
    host_kw = [batch_kwargs(b, pin=True) for b in batches]
    dev_kw = [batch_kwargs(b, device=dev) for b in batches]
    h2d = sum(sum(v.numel() * v.element_size() for v in kw.values() if torch.is_tensor(v)) for kw in host_kw)

    def one_pass(kws, gather):
        outs, n0 = [], gen.kernel_launches
        with torch.no_grad():
            for o in gen.forward_many(kws, lanes=args.lanes, rows=args.rows):
                outs += o
        n_launch = gen.kernel_launches - n0
        if gather and world > 1:
            gather_outputs(list(range(len(outs))), outs, dev)
        return outs, n_launch

    def timed(kws, gather, steps, warmup, sampler=None):
        for _ in range(warmup):
            one_pass(kws, gather)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        if sampler:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        toks = launches = 0
        for _ in range(steps):
            outs, nl = one_pass(kws, gather)
            toks += emitted(outs)
            launches += nl
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = e0.elapsed_time(e1)
        clocks = sampler.stop() if sampler else None
        t = torch.tensor([ms, float(toks), float(launches)], dtype=torch.float64, device=dev)
        if world > 1:
            tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
            return float(tmax[0]), float(tsum[1]), float(tsum[2]), clocks
        return ms, float(toks), float(launches), clocks

    sampler = ClockSampler(local_rank) if rank == 0 else None
    ms, toks, launches, clocks = timed(dev_kw, False, args.steps, args.warmup, sampler)
    ms_e, toks_e, _, _ = timed(host_kw, True, args.steps, 1)
    d2h = int(toks_e / max(args.steps, 1) / world) * 4
    line = {
        "metric": "decoded tokens/sec at beam 4", "value": toks / (ms / 1000.0), "unit": "tokens/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": "configs[1]: MassSeq2Seq base V=60000 6+6 d=768 beam 4, %d synthetic sentences/rank len 16-128 "
                               "length-sorted, --batch %d --capacity %d (%d batches/rank), random-init weights" % (
                                   args.sentences, args.batch, args.capacity, len(batches)),
                   "beam": BEAM, "lanes": args.lanes, "superbatch_rows": args.rows, "sentences_per_rank": args.sentences, "batches_per_rank": len(batches),
                   "l2": "working set > L2: 191 MB of bf16 weights streamed every search step plus the KV cache",
                   "parallelism": "dp%d (whole batches per rank, one all_gather of outputs at the end)" % world},
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
        if args.dtype != "fp32":
            line["roofline"] = time_kernel_rooflines(gen, model, peaks, args.rows)
            line["roofline"]["peak_source"] = "MEASURED_PEAKS.json bf16_tflops (burst, kernel timed alone)" if peaks else "fallback 1590 TFLOP/s"
        if world == 1 and not args.no_cpu_baseline:
            cb = run_reference(argparse.Namespace(**dict(vars(args), steps=1, warmup=0)), 0)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
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
        run_b200(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
