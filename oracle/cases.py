"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Seeded parity cases shared by oracle/make_golden.py
(which runs the REAL reference on them here) and by tests/ (which rebuild the very same
weights and inputs on any box from the seed; torch's CPU generator is deterministic for
the pinned torch build of this image)."""
import hashlib

import torch

from . import restated as R

# name -> recipe.  All models are hidden 768 / 12 heads / FFN 3072 (src/lm_config.py:4-20).
CASES = {
    # text memory, shared enc/dec self-attention (src/seq2seq.py:63-65), ragged source lengths
    "text_b4": dict(kind="text", V=1000, enc=2, dec=2, lang_dec=False, tie=False, B=5, S=14, beam=4, seed=11,
                    eos_bias=0.0),
    # per-language decoders + tie_embed (src/seq2seq.py:67-77), beam 1 => no length penalty (:200-201)
    "text_b1_langdec": dict(kind="text", V=700, enc=1, dec=1, lang_dec=True, tie=True, B=3, S=9, beam=1, seed=12,
                            eos_bias=0.0),
    # EOS is likely: exercises finished-row zeroing, forced PAD, cloning, early termination (:134,195,213)
    "text_eos": dict(kind="text", V=600, enc=1, dec=2, lang_dec=False, tie=False, B=6, S=12, beam=4, seed=13,
                     eos_bias=1.4),
    # image memory (B,49,768) through image_embed= (src/seq_gen.py:83-84,100-103), fixed max_len
    "image_b3": dict(kind="image", V=800, enc=1, dec=2, lang_dec=False, tie=False, B=3, S=49, beam=3, seed=14,
                     eos_bias=0.6, max_len=14),
    # wide beam, (not lang_dec) and tie_embed => single output layer module (src/seq2seq.py:54-59)
    "text_b8_tie": dict(kind="text", V=900, enc=1, dec=1, lang_dec=False, tie=True, B=2, S=10, beam=8, seed=15,
                        eos_bias=0.7),
    # real size (BASELINE configs[1] model: MassSeq2Seq base 6+6, d 768, V 60000), S = 64 ragged; the loop bound is set
    # explicitly so the full-prefix reference finishes in seconds on CPU.  Fixture also holds per-step log-probs.
    "real_6x6": dict(kind="text", V=60000, enc=6, dec=6, lang_dec=False, tie=False, B=4, S=64, beam=4, seed=16,
                     eos_bias=0.0, max_len=20, min_len=40, record=True),
    # src_sizes 90 and 170 at max_len_a = 1.3: the reference computes max_lens in float32 on tensor elements
    # (src/seq_gen.py:113-121), which gives 121 / 225 where double arithmetic gives 122 / 226.  No EOS => output length
    # = max_lens[r], and the over-length zeroing / forced PAD (:138-139,197,207) start at that step.
    "text_maxlens_f32": dict(kind="text", V=500, enc=1, dec=1, lang_dec=False, tie=False, B=3, S=180, beam=2, seed=17,
                             eos_bias=-30.0, lens=[180, 90, 170]),
}

GEN = dict(max_len_a=1.3, max_len_b=5, len_penalty_ratio=0.8)   # src/translate.py:26-28 defaults


def make_text_batch(B, S, V, seed, min_len=4, lens=None):
    """Synthetic padded source batch in MTDataset's format (src/dataset.py:141-163):
    row = <lang 5> body </s 4> pad 0...; pad_idx = first pad position, or S-1 when none."""
    g = torch.Generator().manual_seed(seed)
    if lens is None:
        lens = torch.randint(min_len, S + 1, (B,), generator=g)
        lens[0] = S
    else:
        lens = torch.tensor(lens)
    src = torch.zeros(B, S, dtype=torch.long)
    for b in range(B):
        n = int(lens[b])
        src[b, :n] = torch.randint(7, V, (n,), generator=g)
        src[b, 0] = 5
        src[b, n - 1] = 4
    mask = src != 0
    pad_idx = torch.tensor([S - 1 if int(lens[b]) == S else int(lens[b]) for b in range(B)])
    return src, mask, pad_idx


def build(name):
    """-> (model, call kwargs for beam_search / BeamDecoder.forward, recipe)"""
    c = CASES[name]
    tp = R.FakeTextProcessor(c["V"])
    torch.manual_seed(c["seed"])
    model = R.Seq2SeqModel(tp, lang_dec=c["lang_dec"], tie_embed=c["tie"], enc_layer=c["enc"],
                           dec_layer=c["dec"]).eval()
    B = c["B"]
    tgt = 1
    ol = model.output_layer if ((not c["lang_dec"]) and c["tie"]) else model.output_layer[tgt]
    with torch.no_grad():
        ol.layer.bias[tp.sep_token_id()] += c["eos_bias"]
    kw = dict(first_tokens=torch.full((B,), 6, dtype=torch.long), tgt_langs=torch.full((B,), tgt, dtype=torch.long),
              pad_idx=0, beam_width=c["beam"])
    if c["kind"] == "text":
        src, mask, pad_idx = make_text_batch(B, c["S"], c["V"], c["seed"] + 1000, c.get("min_len", 4), c.get("lens"))
        kw.update(src_inputs=src, src_mask=mask, src_sizes=pad_idx, src_langs=torch.zeros(B, dtype=torch.long))
        if "max_len" in c:
            kw.update(max_len=c["max_len"])
    else:
        g = torch.Generator().manual_seed(c["seed"] + 2000)
        kw.update(image_embed=torch.randn(B, c["S"], 768, generator=g), max_len=c["max_len"])
    return model, kw, c


def weights_digest(model):
    h = hashlib.sha256()
    for k, v in sorted(model.state_dict().items()):
        h.update(k.encode())
        h.update(v.detach().cpu().contiguous().numpy().tobytes())
    return h.hexdigest()
