"""ORACLE -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Tie / near-tie bookkeeping for the 16-bit parity protocol (SURVEY.md §7.2-2, §7.2-3, §8c "tie cases reported").

A 16-bit decode that follows the fp32 reference exactly until two candidates of one sentence are closer than the
arithmetic can resolve, and then picks the other one, is a NEAR-TIE divergence: every later token of that sentence
differs, but no kernel is wrong.  This module finds, per sentence, the first search step at which the product's beam
state leaves the oracle's, and measures the gap between the two choices UNDER THE ORACLE'S OWN fp32 SCORING
(src/seq_gen.py:199-203: candidate score = (beam score + log-prob) / length penalty).
"""
import torch

from . import restated as R


def _candidates(prev, cur, eos, beam, alpha, max_lens):
    """Oracle candidate scores (B, beam_prev * V) of the step that produced trace entry `cur` from state `prev`."""
    i = cur["i"]
    logp = cur["log_probs"].clone()
    B = cur["prefixes"].size(0)
    if prev is None:
        return logp.view(B, -1)                                          # step 1: scores 0, sizes 0 => penalty 1
    pre = prev["prefixes"].view(-1, prev["prefixes"].size(-1))
    fin = torch.any(pre == eos, 1)
    lim = (max_lens < (i + 1)).unsqueeze(-1).expand(-1, beam).reshape(-1)
    logp[fin] = 0
    logp[lim.to(logp.device)] = 0
    s = prev["scores"].view(-1, 1)
    if beam > 1:
        return ((s + logp) / R.len_penalty(prev["sizes"].view(-1), alpha)).view(B, -1)
    return (s + logp).view(B, -1)


def first_divergence(oracle_trace, product_prefixes, beam, V, eos, alpha, max_lens):
    """oracle_trace: list from restated.beam_search(trace=...).  product_prefixes: list over steps of (B*beam, i+1) int
    tensors (the product's beam state AFTER step i; engine trace key "tokens").
    Returns {sentence: dict(step, gap, product_idx, oracle_idx)} for every sentence whose beam state leaves the oracle's;
    gap = max over beam slots of |oracle score of the product's j-th choice - oracle score of the oracle's j-th choice|."""
    out = {}
    B = oracle_trace[0]["prefixes"].size(0)
    max_lens = max_lens.cpu()
    for k in range(min(len(oracle_trace), len(product_prefixes))):
        cur = oracle_trace[k]
        want = cur["prefixes"].cpu()                                     # (B, beam, i+1)
        got = product_prefixes[k].cpu().long().view(B, beam, -1)
        prev = oracle_trace[k - 1] if k > 0 else None
        cand = None
        for b in range(B):
            if b in out or torch.equal(want[b], got[b]):
                continue
            if cand is None:
                cand = _candidates(prev, cur, eos, beam, alpha, max_lens).cpu()
            old = prev["prefixes"].cpu()[b] if prev is not None else want[b][:1, :-1]      # (beam_prev, i)
            idx_p, idx_o = [], []
            for j in range(beam):
                for which, row in ((idx_p, got[b, j]), (idx_o, want[b, j])):
                    src = [s for s in range(old.size(0)) if torch.equal(old[s], row[:-1])]
                    which.append((src[0] if src else 0) * V + int(row[-1]))
            gaps = [abs(float(cand[b, p]) - float(cand[b, o])) for p, o in zip(idx_p, idx_o)]
            out[b] = dict(step=cur["i"], gap=max(gaps), product_idx=idx_p, oracle_idx=idx_o)
    return out


def agreement(got, want):
    """Token agreement over whole output rows (position by position, length mismatch counts as disagreement)."""
    agree = total = 0
    for g, w in zip(got, want):
        g, w = g.tolist(), w.tolist()
        total += max(len(g), len(w))
        agree += sum(1 for a, b in zip(g, w) if a == b)
    return agree, total
