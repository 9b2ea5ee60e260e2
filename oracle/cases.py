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
    # lexical proposals on the last state (src/seq2seq.py:110-144, src/seq_gen.py:191-192): ragged proposal lists padded
    # with PAD (which the reference attends like any other id), one sentence with only PAD proposals (context := 1e-8)
    "text_proposals": dict(kind="text", V=700, enc=1, dec=1, lang_dec=False, tie=False, B=4, S=10, beam=3, seed=18,
                           eos_bias=0.8, proposals=5),
    # object-feature stream (src/seq_gen.py:167-179): a second decoder (`obj_decoder`, src/image_model.py:279-296) over a
    # second memory, mixed by sigmoid(multistream_attention_gate); both memories handed over by a patched encode(images=)
    "image_obj_stream": dict(kind="image_obj", V=650, enc=1, dec=2, lang_dec=False, tie=False, B=3, S=49, S2=7, beam=4, seed=19,
                             eos_bias=0.0, max_len=12),
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


class _PatchedEncode(object):
    """Stands in for ImageCaptioning.encode(images=) (src/image_model.py:298-306): hands out fixed (grid memory, object
    memory) tensors, so that both sides of a parity run decode the same memories."""

    def __init__(self, memory, obj):
        self.memory, self.obj = memory, obj

    def __call__(self, src_inputs=None, src_mask=None, src_langs=None, images=None):
        dev = images.device
        return self.memory.to(dev), self.obj.to(dev)


def attach_obj_stream(model, decoder_cls, dec_config, seed):
    """The attributes BeamDecoder.forward reads for the object stream (src/seq_gen.py:168-177), as
    ImageCaptioning.__init__ creates them (src/image_model.py:279-296); gate values drawn so that the mix is not 50/50."""
    torch.manual_seed(seed)
    model.obj_decoder = decoder_cls(dec_config)
    model.multistream_attention_gate = torch.nn.Parameter(torch.randn(1, dec_config.hidden_size))
    return model


def build(name):
    """-> (model, call kwargs for BeamDecoder.forward, recipe)"""
    c = CASES[name]
    tp = R.FakeTextProcessor(c["V"])
    torch.manual_seed(c["seed"])
    model = R.Seq2SeqModel(tp, lang_dec=c["lang_dec"], tie_embed=c["tie"], enc_layer=c["enc"],
                           dec_layer=c["dec"], use_proposals="proposals" in c)
    if "proposals" in c:
        with torch.no_grad():
            model.lexical_gate.normal_()
    if c["kind"] == "image_obj":
        attach_obj_stream(model, R.DecoderModel, model.decoder.config, c["seed"] + 500)
    model = model.eval()
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
        if "proposals" in c:
            g = torch.Generator().manual_seed(c["seed"] + 3000)
            prop = torch.randint(7, c["V"], (B, c["proposals"]), generator=g)
            prop[1, 3:] = 0
            prop[2, :] = 0                                     # only PAD proposals
            prop[3, 1:] = 0
            kw.update(proposals=prop)
    else:
        g = torch.Generator().manual_seed(c["seed"] + 2000)
        mem = torch.randn(B, c["S"], 768, generator=g)
        if c["kind"] == "image":
            kw.update(image_embed=mem, max_len=c["max_len"])
        else:
            obj = torch.randn(B, c["S2"], 768, generator=g)
            obj[1, 4:] = 0                                     # images with fewer objects: zero rows, still attended (no mask)
            obj[2, 2:] = 0
            model.encode = _PatchedEncode(mem, obj)
            kw.update(images=torch.zeros(B, 1), max_len=c["max_len"])
    return model, kw, c


def run_oracle(model, kw, **opts):
    """The restated oracle on the keyword arguments of BeamDecoder.forward (maps a patched encode(images=) to the two
    memories of the object stream)."""
    kw = dict(kw)
    if "images" in kw:
        mem, obj = model.encode(images=kw.pop("images"))
        kw["image_embed"] = mem
        if obj is not None:
            opts.update(memory2=obj, decoder2=model.obj_decoder, gate=model.multistream_attention_gate)
    return R.beam_search(model, **GEN, **kw, **opts)


def weights_digest(model):
    h = hashlib.sha256()
    for k, v in sorted(model.state_dict().items()):
        h.update(k.encode())
        h.update(v.detach().cpu().contiguous().numpy().tobytes())
    return h.hexdigest()
