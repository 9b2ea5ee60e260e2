"""Decode variants behind the same BeamDecoder API (-m gpu), SURVEY.md §8(f)-3:
  * object-feature stream: second decoder + sigmoid gate (reference src/seq_gen.py:167-179)   -- fixture from the REAL reference
  * lexical proposals: attend_proposal on the last state (src/seq2seq.py:110-144, seq_gen.py:191-192) -- fixture from the REAL reference
  * text + image: the same decoder run over both memories + multimodal gate (src/seq_gen.py:180-188)  -- against the restated oracle
    (the reference's own branch cannot execute, see test docstring)
  * the chained image_embed= calls of src/translate_img.py:53-93
plus the two small kernels on their own against torch.
"""
import ctypes as C

import pytest
import torch
import torch.nn as nn
from torch.nn.utils.rnn import pad_sequence

from imagetranslate_b200 import BeamDecoder, _lib
from oracle import cases, restated as R
from gpu_util import DEV, ACT, stream

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [_lib.F32, _lib.F16, _lib.BF16])
def test_gate_mix_kernel(dtype):
    M, d = 37, 768
    torch.manual_seed(3)
    a, b, gate = torch.randn(M, d, device=DEV), torch.randn(M, d, device=DEV), torch.randn(d, device=DEV)
    x, first = b.clone(), torch.full((M, d), 7.0, device=DEV)
    x_act = torch.zeros(M, d, dtype=ACT[dtype], device=DEV)
    l = _lib.lib()
    _lib.check(l.itb_save_state(None, _lib.ptr(a), _lib.ptr(first), M, d, stream()), "save_state")
    _lib.check(l.itb_gate_mix(dtype, None, _lib.ptr(first), _lib.ptr(gate), 1e-7, _lib.ptr(x), _lib.ptr(x_act), M, d, stream()), "gate_mix")
    torch.cuda.synchronize()
    sig = torch.sigmoid(gate + 1e-7)
    want = sig * a + (1 - sig) * b
    assert torch.equal(first, a)
    assert float((x - want).abs().max()) < 2e-6
    assert float((x_act.float() - want).abs().max()) < (2e-6 if dtype == _lib.F32 else 3e-2)


@pytest.mark.parametrize("beam,P", [(1, 1), (4, 5), (3, 40)])
def test_attend_proposal_kernel_matches_the_reference_arithmetic(beam, P):
    B, d, V = 5, 768, 300
    M = B * beam
    torch.manual_seed(5)
    tp = R.FakeTextProcessor(V)
    model = R.Seq2SeqModel(tp, lang_dec=False, enc_layer=1, dec_layer=1, use_proposals=True).eval()
    with torch.no_grad():
        model.lexical_gate.normal_()
        model.lexical_layer_norm.weight.normal_(1.0, 0.2)
        model.lexical_layer_norm.bias.normal_(0.0, 0.2)
    prop = torch.randint(7, V, (B, P))
    prop[1, P // 2:] = 0
    prop[2, :] = 0
    h = torch.randn(M, d)
    with torch.no_grad():
        want = model.attend_proposal(h.clone(), prop, 0)
    x = h.to(DEV).clone()
    x_act = torch.zeros(M, d, device=DEV)
    f = lambda t: t.detach().to(DEV).float().contiguous()
    emb, gate = f(model.proposal_embedding.weight), f(model.lexical_gate).reshape(-1)
    g, b = f(model.lexical_layer_norm.weight), f(model.lexical_layer_norm.bias)
    pr = prop.to(DEV).to(torch.int32).contiguous()
    _lib.check(_lib.lib().itb_attend_proposal(_lib.F32, None, _lib.ptr(pr), P, 0, _lib.ptr(emb), _lib.ptr(gate), _lib.ptr(g), _lib.ptr(b),
                                              float(model.lexical_layer_norm.eps), _lib.ptr(x), _lib.ptr(x_act), M, beam, d, stream()),
               "attend_proposal")
    torch.cuda.synchronize()
    assert float((x.cpu() - want).abs().max()) < 2e-5
    assert torch.equal(x, x_act)


def _text_image_model(seed, V, dec):
    tp = R.FakeTextProcessor(V)
    torch.manual_seed(seed)
    model = R.Seq2SeqModel(tp, lang_dec=False, tie_embed=False, enc_layer=1, dec_layer=dec)
    model.multimodal_attention_gate = nn.Parameter(torch.randn(1, 768))            # src/image_model.py:139-140
    return model.eval()


class _GridFeatures(nn.Module):
    """Stands in for ModifiedResnet (src/image_model.py:24-82): returns the (grid memory, object features) pair."""

    def __init__(self, mem):
        super().__init__()
        self.register_buffer("mem", mem)

    def forward(self, images):
        return self.mem, None


def test_text_plus_image_streams_share_one_decoder():
    """src/seq_gen.py:105-106,180-188: text memory and image memory through the SAME decoder, mixed by
    sigmoid(multimodal_attention_gate + 1e-7).  The reference's own branch cannot execute as written: ImageMassSeq2Seq.encode
    (src/image_model.py:145-155) returns the image model's (grid, object) TUPLE as `image_embeddings`, which
    BertDecoderModel.forward then calls `.size()` on (src/bert_seq2seq.py:73); `src_langs` stays 1-D (:93-95 expands it only in
    the text-only branch) and the image memory is never repeated per beam.  What the branch means is fixed by the training
    forward of the same model (src/image_model.py:205-216); the restated oracle implements that, and the CUDA path must match
    it token for token in fp32."""
    B, S, V, beam = 4, 11, 900, 4
    model = _text_image_model(71, V, 2)
    src, mask, pad_idx = cases.make_text_batch(B, S, V, 72)
    g = torch.Generator().manual_seed(73)
    mem2 = torch.randn(B, 49, 768, generator=g)
    model.image_model = _GridFeatures(mem2)
    kw = dict(src_inputs=src, src_mask=mask, src_sizes=pad_idx, src_langs=torch.zeros(B, dtype=torch.long),
              first_tokens=torch.full((B,), 6), tgt_langs=torch.ones(B, dtype=torch.long), pad_idx=0, max_len=14)
    want = R.beam_search(model, beam_width=beam, memory2=mem2, decoder2=model.decoder, gate=model.multimodal_attention_gate,
                         **cases.GEN, **kw)
    text_only = R.beam_search(model, beam_width=beam, **cases.GEN, **kw)
    assert [w.tolist() for w in want] != [t.tolist() for t in text_only]           # the second stream matters
    bd = BeamDecoder(model.to(DEV), beam_width=beam, dtype="fp32", **cases.GEN)
    with torch.no_grad():
        got = bd(images=torch.zeros(B, 3, 8, 8), **kw)
    assert [g_.tolist() for g_ in got] == [w.tolist() for w in want]
    # 16-bit production arithmetic runs the same chain (two KV caches, gate mix): sane output of the same shape
    bd16 = BeamDecoder(model, beam_width=beam, dtype="fp16", **cases.GEN)
    with torch.no_grad():
        got16 = bd16(images=torch.zeros(B, 3, 8, 8), **kw)
    same = sum(int(a.tolist() == b.tolist()) for a, b in zip(got16, want))
    print("text+image fp16: %d/%d sentences identical to the fp32 oracle" % (same, B))
    assert same >= B - 1


def test_translate_img_chain_of_image_embed_calls():
    """src/translate_img.py:53-93: three chained generator calls, each decoding from an image_embed (B,49,d) that a
    text-to-image model produced from the previous call's padded outputs, with a caller-computed max_len."""
    B, S, V, beam = 5, 9, 800, 4
    tp = R.FakeTextProcessor(V)
    torch.manual_seed(81)
    model = R.Seq2SeqModel(tp, lang_dec=False, tie_embed=False, enc_layer=1, dec_layer=2).eval()
    with torch.no_grad():
        model.output_layer[0].layer.bias[4] += 0.6
        model.output_layer[1].layer.bias[4] += 0.6
    proj = torch.randn(768, 49 * 768, generator=torch.Generator().manual_seed(82)) * 0.02

    def txt2img(ids, mask, langs):           # stand-in for Caption2Image.forward (src/image_model.py:405-434): (B, 49*d)
        e = model.encoder.embeddings.word_embeddings.weight.detach().cpu()[ids.cpu()] * mask.cpu().unsqueeze(-1)
        return (e.sum(1) / mask.cpu().sum(1, keepdim=True).clamp(min=1)) @ proj

    src, mask, _ = cases.make_text_batch(B, S, V, 83)
    src_langs, dst_langs = torch.zeros(B, dtype=torch.long), torch.ones(B, dtype=torch.long)
    first_dst, first_src = torch.full((B,), 6), src[:, 0]
    max_len = min(int(cases.GEN["max_len_a"] * src.size(1) + cases.GEN["max_len_b"]), 512)

    def chain(generate):
        outs = []
        emb = txt2img(src, mask, src_langs).view(B, 49, -1)
        o1 = generate(first_tokens=first_dst, max_len=max_len, tgt_langs=dst_langs, image_embed=emb, pad_idx=0)
        p1 = pad_sequence(o1, batch_first=True, padding_value=0)
        emb2 = txt2img(p1, p1 != 0, dst_langs).view(B, 49, -1)
        o2 = generate(first_tokens=first_src, max_len=max_len, tgt_langs=src_langs, image_embed=emb2, pad_idx=0)
        p2 = pad_sequence(o2, batch_first=True, padding_value=0)
        emb3 = txt2img(p2, p2 != 0, src_langs).view(B, 49, -1)
        o3 = generate(first_tokens=first_dst, max_len=max_len, tgt_langs=dst_langs, image_embed=emb3, pad_idx=0)
        return [[t.tolist() for t in o] for o in (o1, o2, o3)]

    want = chain(lambda **kw: R.beam_search(model, beam_width=beam, **cases.GEN, **kw))
    bd = BeamDecoder(model.to(DEV), beam_width=beam, dtype="fp32", **cases.GEN)
    with torch.no_grad():
        got = chain(lambda **kw: bd(**kw))
    assert got == want


def test_forward_many_packs_the_variants():
    """forward_many with variant batches: proposal batches of different widths must not share a super-batch (PAD proposals
    are attended, so padding a narrower list would change its sentences), object-stream batches may (both memories travel
    per group); every batch must come out as its own forward call gives it."""
    V, beam = 900, 3
    tp = R.FakeTextProcessor(V)
    torch.manual_seed(91)
    model = R.Seq2SeqModel(tp, lang_dec=False, tie_embed=False, enc_layer=1, dec_layer=2, use_proposals=True).eval()
    with torch.no_grad():
        model.lexical_gate.normal_()
        model.output_layer[1].layer.bias[4] += 0.8
    model = model.to(DEV)
    batches = []
    for i, (B, S, P) in enumerate([(4, 9, 5), (3, 10, 5), (5, 11, 3), (2, 12, 3), (6, 9, 7)]):
        src, mask, pad_idx = cases.make_text_batch(B, S, V, 400 + i)
        prop = torch.randint(7, V, (B, P), generator=torch.Generator().manual_seed(500 + i))
        prop[0, P // 2:] = 0
        batches.append(dict(src_inputs=src, src_mask=mask, src_sizes=pad_idx, src_langs=torch.zeros(B, dtype=torch.long),
                            first_tokens=torch.full((B,), 6), tgt_langs=torch.ones(B, dtype=torch.long), pad_idx=0, proposals=prop))
    bd = BeamDecoder(model, beam_width=beam, dtype="fp32", **cases.GEN)
    assert [len(p) for p in bd._pack(batches, 4096)] == [2, 2, 1]
    with torch.no_grad():
        seq = [bd(**kw) for kw in batches]
        par = bd.forward_many(batches, lanes=2, rows=4096)
    assert [[o.tolist() for o in b] for b in par] == [[o.tolist() for o in b] for b in seq]
    cpu = model.cpu()
    want = R.beam_search(cpu, beam_width=beam, **cases.GEN, **batches[2])
    assert [o.tolist() for o in seq[2]] == [w.tolist() for w in want]
    # object stream: two batches with different object counts in one super-batch
    m2, kw_a, c = cases.build("image_obj_stream")
    kw_a = dict(kw_a); kw_a.pop("beam_width")
    mem, obj = m2.encode(images=kw_a["images"])
    kw_b = dict(kw_a, first_tokens=kw_a["first_tokens"][:2], tgt_langs=kw_a["tgt_langs"][:2], images=torch.ones(2, 1))

    class TwoBatches(object):          # hands out the memories of batch a (zeros image) or b (ones image: fewer objects)
        def __call__(self, src_inputs=None, src_mask=None, src_langs=None, images=None):
            if float(images.flatten()[0]) == 0.0:
                return mem.to(images.device), obj.to(images.device)
            return mem[:2].flip(0).contiguous().to(images.device), obj[:2, :4].contiguous().to(images.device)

    m2.encode = TwoBatches()
    m2 = m2.to(DEV)
    bd2 = BeamDecoder(m2, beam_width=4, dtype="fp32", **cases.GEN)
    with torch.no_grad():
        seq2 = [bd2(**kw_a), bd2(**kw_b)]
        par2 = bd2.forward_many([kw_a, kw_b], lanes=1, rows=4096)
    assert [len(p) for p in bd2._pack([kw_a, kw_b], 4096)] == [2]
    assert [[o.tolist() for o in b] for b in par2] == [[o.tolist() for o in b] for b in seq2]
