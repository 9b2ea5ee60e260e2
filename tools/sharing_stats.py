"""How much of the self-attention KV cache the beam slots of a sentence share in the headline workload (not a test, not the
bench): decodes one super-batch of configs[1] and reads the final ancestor table.  For every sentence and position p it
counts the distinct physical rows among the beam slots; reports the bytes the kernel reads (common prefix once + every
slot's tail) against the ideal (every distinct row once) and against no sharing at all.
Usage: python tools/sharing_stats.py [first_batch=40] [n_batches=12]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from imagetranslate_b200 import BeamDecoder  # noqa: E402
from imagetranslate_b200.models import MassSeq2Seq, SyntheticTextProcessor  # noqa: E402

first = int(sys.argv[1]) if len(sys.argv) > 1 else 40
nb = int(sys.argv[2]) if len(sys.argv) > 2 else 12
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = MassSeq2Seq(SyntheticTextProcessor(60000), lang_dec=False, tie_embed=False, enc_layer=6, dec_layer=6).eval().to(dev).to(torch.bfloat16)
gen = BeamDecoder(model, beam_width=4, dtype="fp16", **bench.GEN)
batches = bench.build_batches(4096, 1234, 4000, 600)[first:first + nb]
kws = [bench.batch_kwargs(b, device=dev) for b in batches]
with torch.no_grad():
    outs = gen.forward_many(kws, lanes=1, rows=4096)
eng = list(gen._engines.values())[0]
ws = eng.ws.buf
B = sum(int(k["src_inputs"].shape[0]) for k in kws)
M = B * 4
steps = eng.last_steps
ctl = ws["ctl"][:8].cpu().tolist()
T_max = ws["anc"].numel() // (2 * M) if False else None
# the engine sized anc as [2][M][T_max] for this call: recover T_max from the tokens buffer of the same call
T_max = -(-max(len(o) for b in outs for o in b) // 16) * 16
anc = ws["anc"][:2 * M * T_max].view(2, M, T_max)[ctl[3]].view(B, 4, T_max).long().cpu()
print("super-batch: %d sentences, %d rows, %d steps, T_max %d" % (B, M, steps, T_max))
for t in (steps // 4, steps // 2, steps - 1):
    a = anc[:, :, :t]                                              # (B, 4, t) physical slot of every position
    same = (a == a[:, :1]).all(1)                                  # (B, t): all four slots read the same row
    c = torch.where(same.all(1), torch.full((B,), t), (~same).float().argmax(1))      # common prefix per sentence
    distinct = torch.zeros(B, t)
    for s in range(4):
        distinct += (a == s).any(1).float()
    read = c.float() + 4.0 * (t - c.float())
    print("t = %3d: common prefix %.1f positions on average (%.0f %% of t); rows read per sentence %.1f, ideal (distinct rows) %.1f, "
          "no sharing %d  -> kernel reads %.2fx the ideal, %.2fx less than without sharing" % (
              t, c.float().mean(), 100 * c.float().mean() / t, read.mean(), distinct.sum(1).mean(), 4 * t,
              read.mean() / distinct.sum(1).mean(), 4 * t / read.mean()))
