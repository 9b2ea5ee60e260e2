"""Pins the oracle restatement against the reference's own code, executed where it lies under
/root/reference (only present in the build container; skipped on the GPU box)."""
import pytest
import torch

from oracle import cases, reference_loader, restated as R

pytestmark = pytest.mark.skipif(not reference_loader.available(), reason="/root/reference not present")


def _small(seed, lang_dec, tie, beam, eos_bias, B=4, S=11, V=300):
    tp = R.FakeTextProcessor(V)
    torch.manual_seed(seed)
    m = R.Seq2SeqModel(tp, lang_dec=lang_dec, tie_embed=tie, enc_layer=2, dec_layer=2, embed_dim=96,
                       intermediate_dim=160).eval()
    ol = m.output_layer if ((not lang_dec) and tie) else m.output_layer[1]
    with torch.no_grad():
        ol.layer.bias[4] += eos_bias
    src, mask, pad_idx = cases.make_text_batch(B, S, V, seed + 5)
    kw = dict(src_inputs=src, src_mask=mask, src_sizes=pad_idx, src_langs=torch.zeros(B, dtype=torch.long),
              first_tokens=torch.full((B,), 6), tgt_langs=torch.ones(B, dtype=torch.long), pad_idx=0)
    return m, kw


@pytest.mark.parametrize("lang_dec,tie,beam,eos_bias", [(False, False, 4, 0.0), (True, False, 3, 1.0),
                                                         (False, True, 1, 0.5), (True, True, 5, 1.3),
                                                         (False, False, 2, 2.0)])
def test_restated_equals_reference(lang_dec, tie, beam, eos_bias):
    ns = reference_loader.load()
    m, kw = _small(3, lang_dec, tie, beam, eos_bias)
    ref = ns.mass_seq2seq.MassSeq2Seq(m.text_processor, lang_dec=lang_dec, tie_embed=tie, enc_layer=2, dec_layer=2,
                                      embed_dim=96, intermediate_dim=160).eval()
    assert sorted(ref.state_dict().keys()) == sorted(m.state_dict().keys())
    ref.load_state_dict(m.state_dict())
    bd = ns.seq_gen.BeamDecoder(ref, beam_width=beam, **cases.GEN)
    for unpad in (True, False):
        with torch.no_grad():
            a = bd(unpad_output=unpad, **kw)
        b = R.beam_search(m, beam_width=beam, unpad_output=unpad, **cases.GEN, **kw)
        assert [x.tolist() for x in a] == [x.tolist() for x in b]


def test_reference_shape_test_replayed():
    """src/tests/test_model.py:53-74 only checks the teacher-forced output SHAPE; replay it on the shim."""
    ns = reference_loader.load()
    tp = R.FakeTextProcessor(120)
    model = ns.seq2seq.Seq2Seq(tp, enc_layer=1, dec_layer=1, embed_dim=96, intermediate_dim=128)
    src = torch.tensor([[5, 9, 10, 4, 0], [5, 11, 12, 13, 4]])
    out = model(src_inputs=src, tgt_inputs=src, src_mask=src != 0, tgt_mask=src != 0,
                src_langs=torch.zeros(2, dtype=torch.long), tgt_langs=torch.zeros(2, dtype=torch.long), log_softmax=True)
    assert out.size(-1) == 120
