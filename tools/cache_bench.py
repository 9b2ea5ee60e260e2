"""Decoder-cache behaviour on ragged batches (not a test, not the bench): N reference batches of random shape (B sentences x S
source tokens) through BeamDecoder.forward on the full-size model; reports how many decoders (TMA descriptors + captured
8-step graph) the shape lattice (S rounded up to 8, cache capacity to 16) makes the engine build, and what building them
costs (first pass vs a second pass over the same batches, all cached).  Usage: python tools/cache_bench.py [N=300] [max_len=24]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from imagetranslate_b200 import BeamDecoder  # noqa: E402
from imagetranslate_b200.models import MassSeq2Seq, SyntheticTextProcessor  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 300
MAXLEN = int(sys.argv[2]) if len(sys.argv) > 2 else 24
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = MassSeq2Seq(SyntheticTextProcessor(60000), lang_dec=False, tie_embed=False, enc_layer=6, dec_layer=6).eval().to(dev).to(torch.bfloat16)
gen = BeamDecoder(model, beam_width=4, max_len_a=1.3, max_len_b=5, len_penalty_ratio=0.8, dtype="fp16")
g = torch.Generator().manual_seed(7)
batches = []
for _ in range(N):
    S = int(torch.randint(8, 129, (1,), generator=g))
    B = max(1, min(int(4000 // (S + 1)), int(torch.randint(1, 240, (1,), generator=g))))
    src = torch.randint(7, 60000, (B, S), generator=g)
    src[:, 0], src[:, -1] = 5, 4
    batches.append(dict(src_inputs=src, src_sizes=torch.full((B,), S - 1), first_tokens=torch.full((B,), 6), src_mask=torch.ones(B, S, dtype=torch.bool),
                        src_langs=torch.zeros(B, dtype=torch.long), tgt_langs=torch.ones(B, dtype=torch.long), pad_idx=0, max_len=MAXLEN))
shapes = len(set((b["src_inputs"].shape[0], b["src_inputs"].shape[1]) for b in batches))
with torch.no_grad():
    gen(**batches[0])                          # weights packed, workspace sized for the first shape
    times = []
    for rep in range(2):
        torch.cuda.synchronize()
        t0 = time.time()
        for kw in batches:
            gen(**kw)
        torch.cuda.synchronize()
        times.append(time.time() - t0)
        eng = list(gen._engines.values())[0]
        print("pass %d: %.2f s for %d batches (%d distinct (B,S) shapes), decoders created so far %d, cached %d" % (
            rep, times[-1], N, shapes, eng.decoders_created, len(eng._decoders)))
print("cost of building decoders: %.1f ms per new shape (%.1f %% of the first pass at max_len %d)" % (
    1000 * (times[0] - times[1]) / max(eng.decoders_created - 1, 1), 100 * (times[0] - times[1]) / times[0], MAXLEN))
