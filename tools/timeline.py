"""Launch-by-launch timeline of the REAL search step (graph replay, PDL chain) from the diagnostic build of the library
(imagetranslate_b200/csrc/build_timeline.sh; every kernel stamps %globaltimer at entry and when its programmatic dependency
resolves).  Usage:  ITB_LIB=imagetranslate_b200/csrc/build/libitb_timeline.so python tools/timeline.py B S [dtype] [beam]
Prints, per position of the step's launch chain, the mean time from this kernel's dependency-resolved stamp to the next
kernel's (= what the launch costs the chain) and how long before that the kernel had been resident (entry -> resolved)."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from imagetranslate_b200 import BeamDecoder, _lib  # noqa: E402
from imagetranslate_b200.models import MassSeq2Seq, SyntheticTextProcessor  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 44
S = int(sys.argv[2]) if len(sys.argv) > 2 else 89
dtype = sys.argv[3] if len(sys.argv) > 3 else "fp16"
beam = int(sys.argv[4]) if len(sys.argv) > 4 else 4
FILES = {1: "misc", 2: "attention", 3: "attention_mma", 4: "attention_bulk", 5: "beam", 6: "gemm_simt", 7: "gemm_tc", 8: "api"}
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = MassSeq2Seq(SyntheticTextProcessor(60000), lang_dec=False, tie_embed=False, enc_layer=1, dec_layer=6).eval().to(dev)
model = model.to(torch.bfloat16)
gen = BeamDecoder(model, beam_width=beam, max_len_a=1.3, max_len_b=5, len_penalty_ratio=0.8, dtype=dtype)
code = gen._dtype_code(model.encoder.embeddings.word_embeddings.weight)
eng = gen._engine(dev, code)
pw = gen._weights(model, 1, code, dev)
max_len = min(int(1.3 * S + 5), 512)
group = dict(memory=torch.randn(B, S, 768, device=dev).to(torch.bfloat16), src_mask=torch.ones(B, S, dtype=torch.bool, device=dev),
             first_tokens=torch.full((B,), 6), tgt_type_ids=torch.ones(B, dtype=torch.long), max_lens=torch.full((B,), max_len),
             max_len=max_len)
lib = C.CDLL(_lib.LIB_PATH)
lib.itb_timeline_read.argtypes = [C.c_void_p, C.c_int]
for it in range(2):
    eng.decode_groups(pw, [group], beam, 0, 4, 0.8)
    torch.cuda.synchronize()
assert lib.itb_timeline_reset() == 0
eng.decode_groups(pw, [group], beam, 0, 4, 0.8)
torch.cuda.synchronize()
CAP = 16384
buf = np.zeros(1 + 2 * CAP, dtype=np.uint64)
n = lib.itb_timeline_read(buf.ctypes.data, CAP)
assert n > 0, "no stamps: is ITB_LIB the timeline build?"
tags, times = buf[1:1 + 2 * n:2].astype(np.int64), buf[2:2 + 2 * n:2].astype(np.int64)
order = np.argsort(times, kind="stable")
tags, times = tags[order], times[order]
sub = (tags % 100000) >= 90000                       # ITB_TL(n) stamps inside a kernel
dep = ((tags % 100000) >= 50000) & ~sub
dt, dtime = tags[dep] - 50000, times[dep]
et, etime = tags[~dep & ~sub], times[~dep & ~sub]
st_, stime = tags[sub], times[sub]


def name(tag):
    return "%s.cu:%d" % (FILES.get(int(tag) // 100000, "?"), int(tag) % 100000)


# step boundary = the embedding kernel's dependency stamp (first kernel of a step): the tag that occurs once per period
embed = [t for t in np.unique(dt) if name(t).startswith("misc")]
first = None
for t in embed:
    idx = np.nonzero(dt == t)[0]
    if len(idx) > 4:
        gaps = np.diff(idx)
        if first is None or gaps.min() > first[1]:
            first = (t, gaps.min(), idx)
t0, period, starts = first
print("stamps %d (dependency %d, entry %d); step = %d launches, starting at %s; %d whole steps recorded" % (
    n, dep.sum(), (~dep & ~sub).sum(), period, name(t0), len(starts) - 1))
starts = [s for s in starts[1:-1] if s + period < len(dt)]      # skip the first step (cold) and the last (clipped)
chain = np.array([[dtime[s + k + 1] - dtime[s + k] for k in range(period)] for s in starts], dtype=np.float64) / 1000.0
# entry stamp of the same kernel instance: the latest entry stamp with that tag at or before the dependency stamp
res = np.zeros((len(starts), period))
for i, s in enumerate(starts):
    for k in range(period):
        j = np.searchsorted(etime, dtime[s + k], side="right") - 1
        while j >= 0 and et[j] // 100000 != dt[s + k] // 100000:      # same source file (entry and dependency stamps differ in line)
            j -= 1
        res[i, k] = (dtime[s + k] - etime[j]) / 1000.0 if j >= 0 else float("nan")
print("%-4s %-24s %10s %10s %12s" % ("pos", "kernel (dependency stamp)", "chain us", "min us", "resident us"))
for k in range(period):
    print("%-4d %-24s %10.2f %10.2f %12.2f" % (k, name(dt[starts[0] + k]), chain[:, k].mean(), chain[:, k].min(), np.nanmean(res[:, k])))
print("sum of the chain: %.1f us per step (mean over %d steps)" % (chain.sum(1).mean(), len(starts)))
by = {}
for k in range(period):
    by.setdefault(name(dt[starts[0] + k]), []).append(chain[:, k].mean())
for kname, v in sorted(by.items(), key=lambda kv: -sum(kv[1])):
    print("  %-24s x%-3d %8.1f us per step  (%.2f us per launch)" % (kname, len(v), sum(v), sum(v) / len(v)))
if len(st_):
    print("stamps inside kernels: mean time since the kernel's dependency stamp")
    for t in np.unique(st_):
        d = []
        for tm in stime[st_ == t]:
            j = np.searchsorted(dtime, tm, side="right") - 1
            while j >= 0 and dt[j] // 100000 != t // 100000:
                j -= 1
            if j >= 0:
                d.append((tm - dtime[j]) / 1000.0)
        print("  %s #%d  n=%d  %8.2f us" % (FILES.get(int(t) // 100000, "?"), int(t) % 100000 - 90000, len(d), float(np.mean(d[2:])) if len(d) > 2 else float("nan")))
