"""Parity of the PRODUCTION 16-bit path (-m gpu): the tcgen05 GEMMs, the tensor-core attention kernel, the fused vocabulary
epilogue and the phase-2-only beam merge -- a different kernel set from the fp32 parity mode -- held to the north_star bar:

  * per-step log-prob within 1e-2 abs of the fp32 reference over the WHOLE vocabulary on the same UN-ROUNDED random-init
    weights (the figure against an oracle whose weights were first rounded to bf16 is printed as a second number);
  * >= 99 % token agreement, with every sentence whose beam state leaves the oracle's traced to its first diverging search
    step and classified under the oracle's own fp32 scoring: a divergence whose two choices score within 2e-2 of each other
    is a near-tie ("tie cases reported"); anything else fails;
  * at real size (6+6 layers, d 768, V 60000, S 64) against per-step log-probs recorded inside the REAL reference's loop
    (tests/golden/real_6x6.npz, oracle/make_golden.py), not only against the restatement.

Production arithmetic = "fp16": 16-bit weights held in fp16 (exact for a bf16 checkpoint), fp16 activations / KV cache, fp32
accumulation, softmax, LayerNorm and scores (what BeamDecoder(dtype="auto") picks for a half / bf16 model and bench.py
runs).  It meets every bar above.  Pure bf16 arithmetic (dtype="bf16") is also built and tested; its 8-bit mantissa at every
GEMM input and in the KV cache measures 1.4e-2 over 60000 entries x 6+6 layers, so it is held to that measured bound and
to the near-tie protocol, not advertised as meeting 1e-2.
"""
import json
import os

import numpy as np
import pytest
import torch

from imagetranslate_b200 import BeamDecoder
from oracle import cases, diverge, restated as R

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DEV = "cuda:0"
NEAR_TIE = 2e-2


def _clone_model(model, rounded=False, sharpen=1.0):
    c = model.config
    m = R.Seq2SeqModel(model.text_processor, lang_dec=model.lang_dec, tie_embed=model.tie_embed, enc_layer=model.enc_layer,
                       dec_layer=model.dec_layer, embed_dim=c.hidden_size, intermediate_dim=c.intermediate_size).eval()
    m.load_state_dict(model.state_dict())
    with torch.no_grad():
        if rounded:
            for p in m.parameters():
                if p.dim() >= 2:
                    p.copy_(p.to(torch.bfloat16).float())
        if sharpen != 1.0:
            for name, p in m.named_parameters():
                if name.startswith("output_layer") and name.endswith("layer.weight"):
                    p.mul_(sharpen)
    return m


def _fresh(seed, V, B, S, beam, enc, dec, eos_bias=0.0, min_len=4):
    tp = R.FakeTextProcessor(V)
    torch.manual_seed(seed)
    model = R.Seq2SeqModel(tp, lang_dec=False, tie_embed=False, enc_layer=enc, dec_layer=dec).eval()
    with torch.no_grad():
        model.output_layer[1].layer.bias[4] += eos_bias
    src, mask, pad_idx = cases.make_text_batch(B, S, V, seed + 7, min_len)
    kw = dict(src_inputs=src, src_mask=mask, src_sizes=pad_idx, src_langs=torch.zeros(B, dtype=torch.long),
              first_tokens=torch.full((B,), 6), tgt_langs=torch.ones(B, dtype=torch.long), pad_idx=0)
    return model, kw


def _product(model, kw, mode, beam, traced=True):
    """Decodes on the CUDA path with the per-step parity trace; the fp32 master copy of the weights stays untouched."""
    gpu = _clone_model(model).to(DEV)
    bd = BeamDecoder(gpu, beam_width=beam, dtype=mode, **cases.GEN)
    bd.trace = [] if traced else None
    with torch.no_grad():
        outs = bd(**kw)
    torch.cuda.synchronize()
    return outs, bd.trace


def _oracle_log_probs(oracle_gpu, memory, mask, prefixes, beam, first_step):
    """fp32 log-softmax rows of the oracle (on the GPU, TF32 off) for the given prefixes: teacher forcing."""
    pre = prefixes.to(DEV).long()
    rep = (lambda x: x) if first_step else (lambda x: x.repeat_interleave(beam, 0))
    st = oracle_gpu.decoder(encoder_states=rep(memory), input_ids=pre, encoder_attention_mask=rep(mask),
                            token_type_ids=torch.ones_like(pre))[:, -1]
    return torch.log_softmax(oracle_gpu.output_layer[1](st), -1)


def _teacher_forced_error(model, kw, mode, beam, every=1):
    """max |log p_product - log p_oracle| over the whole vocabulary on the PRODUCT's own prefixes (so search divergence can
    neither hide nor inflate it), against the oracle on un-rounded weights and on bf16-rounded weights; step-wise top-1
    agreement with the un-rounded oracle."""
    assert not torch.backends.cuda.matmul.allow_tf32
    _, trace = _product(model, kw, mode, beam)
    res = {}
    for tag, rounded in (("unrounded", False), ("rounded", True)):
        orc = _clone_model(model, rounded=rounded).to(DEV)
        src, mask = kw["src_inputs"].to(DEV), kw["src_mask"].to(DEV)
        worst = 0.0
        agree = total = 0
        with torch.no_grad():
            memory = orc.encode(src, mask, kw["src_langs"].to(DEV).unsqueeze(-1).expand(-1, src.size(1)))[0]
            for t in trace[::every]:
                first = t["step"] == 1
                pre = t["prefixes"][::beam] if first else t["prefixes"]
                want = _oracle_log_probs(orc, memory, mask, pre, beam, first)
                got = torch.log_softmax(t["logits"].to(DEV).float(), -1)
                got = got[::beam] if first else got
                worst = max(worst, float((got - want).abs().max()))
                agree += int((got.argmax(-1) == want.argmax(-1)).sum())
                total += got.size(0)
        res[tag] = (worst, agree, total)
        del orc
    return res


# (mode, weights the model holds, bound on max |dlogp| over the whole vocabulary vs the fp32 oracle ON THE SAME WEIGHTS,
#  floor of the step-wise top-1 agreement).  "fp16" is the production arithmetic (BeamDecoder dtype="auto" on a half / bf16
#  model, bench.py): 16-bit weights held in fp16 (exact for bf16-valued weights of this magnitude), fp16 activations and KV
#  cache, fp32 accumulation -- it meets the north_star bar with margin.  Pure bf16 arithmetic (dtype="bf16": 8 mantissa
#  bits at every GEMM input and in the KV cache) measures 1.40e-2 at real size and 9.9e-3 on the small model: its bound
#  here is what bf16 rounding gives, documented in DESIGN.md §5, not the north_star bar.
PROTOCOLS = [("fp16", "fp32_master", 1e-2, 0.99), ("fp16", "bf16_valued", 1e-2, 0.99), ("bf16", "fp32_master", 1.6e-2, 0.98)]


@pytest.mark.parametrize("mode,weights,bound,floor", PROTOCOLS, ids=["%s-%s" % (m, w) for m, w, _, _ in PROTOCOLS])
@pytest.mark.parametrize("size", ["2layer_V6000", "6x6_V60000"])
def test_16bit_whole_vocabulary_log_prob_error_on_unrounded_weights(mode, weights, bound, floor, size):
    if size == "2layer_V6000":
        model, kw = _fresh(41, 6000, 6, 14, 4, 1, 2)
        every = 1
    else:
        model, kw = _fresh(45, 60000, 8, 40, 4, 6, 6, min_len=20)
        kw["max_len"] = 24
        every = 2
    if weights == "bf16_valued":            # a bf16 checkpoint / model.bfloat16(): these ARE the model's weights
        model = _clone_model(model, rounded=True)
    res = _teacher_forced_error(model, kw, mode, 4, every)
    (w_u, a_u, n_u), (w_r, _, _) = res["unrounded"], res["rounded"]
    print("%s %s on %s weights, teacher-forced whole-vocabulary max |dlogp|: %.3e vs the fp32 oracle on the same weights (%.3e vs "
          "an oracle whose weights were rounded to bf16); step-wise top-1 agreement %d/%d = %.4f" % (
              size, mode, weights, w_u, w_r, a_u, n_u, a_u / n_u))
    assert w_u <= bound, w_u
    assert a_u >= floor * n_u


@pytest.mark.parametrize("mode,tol", [("fp32", 2e-4), ("fp16", 1e-2), ("bf16", 1.6e-2)])
def test_real_size_per_step_log_probs_against_the_recorded_reference_run(mode, tol):
    """6+6 layers, V 60000, S 64: log-softmax rows produced by the CUDA path against those recorded inside the REAL
    reference's loop (vocabulary subsample + the 8 largest entries of every row), on every row-step whose prefix the
    product shares with the reference run; fp32 must share all of them (token identity)."""
    z = np.load(os.path.join(GOLD, "real_6x6.npz"))
    meta = json.loads(str(z["meta"]))
    model, kw, c = cases.build("real_6x6")
    assert cases.weights_digest(model) == meta["weights_sha256"]
    kw = dict(kw)
    beam = kw.pop("beam_width")
    outs, trace = _product(model, kw, mode, beam)
    sub = torch.from_numpy(z["sub_index"])
    B = len(outs)
    worst = worst_top = 0.0
    shared = total = top1 = 0
    for k, t in enumerate(trace):
        got = torch.log_softmax(t["logits"].float(), -1)
        if k == 0:
            rows_p, rows_f = list(range(0, B * beam, beam)), list(range(B))
        else:
            before = torch.from_numpy(z["step_prefixes"][k - 1, :, :k + 1])
            rows_p = [r for r in range(B * beam) if torch.equal(t["prefixes"][r].long(), before[r])]
            rows_f = rows_p
        total += B if k == 0 else B * beam
        shared += len(rows_p)
        if not rows_p:
            continue
        g = got[rows_p]
        worst = max(worst, float((g[:, sub] - torch.from_numpy(z["step_logp_sub"][k, rows_f])).abs().max()))
        ti = torch.from_numpy(z["step_top_idx"][k, rows_f])
        worst_top = max(worst_top, float((g.gather(1, ti) - torch.from_numpy(z["step_top_val"][k, rows_f])).abs().max()))
        top1 += int((g.argmax(-1) == ti[:, 0]).sum())
    print("real_6x6 %s vs recorded reference: max |dlogp| subsample %.3e, reference top-8 %.3e; %d/%d row-steps share the "
          "reference prefix, top-1 agreement on them %d/%d" % (mode, worst, worst_top, shared, total, top1, shared))
    assert worst <= tol and worst_top <= min(tol, 1e-2)
    if mode == "fp32":
        assert shared == total
        assert [o.tolist() for o in outs] == [z["tokens"][r, :z["lengths"][r]].tolist() for r in range(B)]
    else:
        assert shared >= 0.5 * total and top1 >= (0.99 if mode == "fp16" else 0.98) * shared
        if mode == "fp16":
            assert shared == total and [o.tolist() for o in outs] == [z["tokens"][r, :z["lengths"][r]].tolist() for r in range(B)]


@pytest.mark.parametrize("mode,sharpen", [("fp16", 1.0), ("fp16", 8.0), ("bf16", 1.0)],
                         ids=["fp16-raw_init", "fp16-sharpened_x8", "bf16-raw_init"])
def test_16bit_free_running_agreement_with_tie_report(mode, sharpen):
    """Free-running search against the fp32 oracle on the same un-rounded weights (raw init, and the sharpened-weights
    protocol of SURVEY §7.2-3: output_layer.*.layer.weight x 8).  Raw token agreement is printed -- and for the production
    arithmetic (fp16) asserted >= 99 % outright; every sentence whose beam state leaves the oracle's (hidden beam slots
    included) is traced to its first diverging step and must be a near-tie there (gap < 2e-2 under the oracle's fp32
    scores).  With the near-tie sentences set aside, token agreement must be >= 99 %.
    Scaling the output layer by 8 scales every log-prob difference AND every arithmetic error by 8 (they all pass through
    that layer), so it cannot turn near-ties into clear decisions; it is kept for fp16, which passes it outright, and not
    run for pure bf16, whose errors (x 8) would be held against an unscaled 2e-2 gap."""
    model, kw = _fresh(51, 8000, 24, 16, 4, 1, 6, min_len=8)
    model = _clone_model(model, sharpen=sharpen)
    beam = 4
    orc = _clone_model(model).to(DEV)
    otrace = []
    want = R.beam_search(orc, beam_width=beam, trace=otrace, **cases.GEN, **kw)
    otrace = [{k: (v.cpu() if torch.is_tensor(v) else v) for k, v in t.items()} for t in otrace]
    got, trace = _product(model, kw, mode, beam)
    mlf = lambda s: min(int(cases.GEN["max_len_a"] * s + cases.GEN["max_len_b"]), 512)
    max_lens = torch.LongTensor([mlf(x) for x in kw["src_sizes"]])
    div = diverge.first_divergence(otrace, [t["tokens"] for t in trace], beam, 8000, 4, cases.GEN["len_penalty_ratio"], max_lens)
    raw_a, raw_n = diverge.agreement(got, want)
    differing = [b for b in range(len(got)) if got[b].tolist() != want[b].tolist()]
    for b in differing:
        assert b in div, "sentence %d differs in its output but never left the oracle's beam state" % b
    ties = {b: d for b, d in div.items() if d["gap"] < NEAR_TIE}
    hard = {b: d for b, d in div.items() if d["gap"] >= NEAR_TIE}
    keep = [b for b in range(len(got)) if b not in ties]
    adj_a, adj_n = diverge.agreement([got[b] for b in keep], [want[b] for b in keep])
    print("%s %s: raw token agreement %d/%d = %.4f; %d/%d sentences differ; beam state diverged for %d sentences, %d near-ties "
          "(max gap %.2e), %d not; agreement without the near-tie sentences %d/%d" % (
              mode, "sharpened x8" if sharpen != 1.0 else "raw init", raw_a, raw_n, raw_a / raw_n, len(differing), len(got),
              len(div), len(ties), max([d["gap"] for d in ties.values()] + [0.0]), len(hard), adj_a, adj_n))
    for b, d in sorted(div.items()):
        print("   sentence %d: first diverging step %d, oracle-score gap %.3e" % (b, d["step"], d["gap"]))
    assert not hard, hard
    assert adj_a >= 0.99 * adj_n
    if mode == "fp16":
        assert raw_a >= 0.99 * raw_n
