"""Average time of one search step of the real decode path (graph replay, PDL chain, split-K ...) at a given shape,
with synthetic memory and random-init weights.  Usage: python tools/step_bench.py B S [groups] [dtype] [beam]
The decode runs the full max_len = int(1.3*S+5) steps; reported: us/step, us/row-step, steps."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from imagetranslate_b200 import BeamDecoder  # noqa: E402
from imagetranslate_b200.models import MassSeq2Seq, SyntheticTextProcessor  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 61
S = int(sys.argv[2]) if len(sys.argv) > 2 else 64
G = int(sys.argv[3]) if len(sys.argv) > 3 else 1
dtype = sys.argv[4] if len(sys.argv) > 4 else "bf16"
beam = int(sys.argv[5]) if len(sys.argv) > 5 else 4
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = MassSeq2Seq(SyntheticTextProcessor(60000), lang_dec=False, tie_embed=False, enc_layer=1, dec_layer=6).eval().to(dev)
if dtype != "fp32":
    model = model.to(torch.bfloat16)
gen = BeamDecoder(model, beam_width=beam, max_len_a=1.3, max_len_b=5, len_penalty_ratio=0.8, dtype=dtype)
code = gen._dtype_code(model.encoder.embeddings.word_embeddings.weight)
eng = gen._engine(dev, code)
pw = gen._weights(model, 1, code, dev)
max_len = min(int(1.3 * S + 5), 512)
groups = []
for g in range(G):
    groups.append(dict(memory=torch.randn(B, S, 768, device=dev).to(model.encoder.embeddings.word_embeddings.weight.dtype),
                       src_mask=torch.ones(B, S, dtype=torch.bool, device=dev), first_tokens=torch.full((B,), 6),
                       tgt_type_ids=torch.ones(B, dtype=torch.long), max_lens=torch.full((B,), max_len), max_len=max_len))
for it in range(3):
    torch.cuda.synchronize()
    t0 = time.time()
    eng.decode_groups(pw, groups, beam, 0, 4, 0.8)
    torch.cuda.synchronize()
    dt = time.time() - t0
steps = eng.last_steps
print("B=%d x %d groups  S=%d  rows=%d  steps=%d : %.1f us/step, %.3f us/row-step, %.0f tok/s (emitted, beam %d)" % (
    B, G, S, B * G * beam, steps, dt * 1e6 / steps, dt * 1e6 / steps / (B * G * beam), B * G * steps / dt, beam))
