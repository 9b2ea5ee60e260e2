"""Anchors of the oracle's third-party arithmetic (CPU, any box).

The reference star-imports `transformers==2.9.0` BERT classes (src/bert_seq2seq.py:3; pinned src/requirements.txt:8), which
cannot be installed offline; oracle/shim/transformers restates them.  Both the real-reference
fixtures and oracle/restated.py run on that shim, so the shim itself is pinned here against an INDEPENDENT implementation
of the same published architecture: the installed transformers' eager `BertModel`, loaded with the shim's own state dict
(identical parameter names) -- encoder with a padding mask, and causal decoder with cross-attention
and a memory padding mask.  Plus the property the incremental (KV-cache) formulation rests on: causal prefix invariance
(SURVEY Appendix C.6).
"""
import pytest
import torch

from oracle import restated as R

transformers = pytest.importorskip("transformers")


def _hf_config(cfg, **extra):
    c = transformers.BertConfig(
        vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size, num_hidden_layers=cfg.num_hidden_layers,
        num_attention_heads=cfg.num_attention_heads, intermediate_size=cfg.intermediate_size, hidden_act="gelu",
        max_position_embeddings=cfg.max_position_embeddings, type_vocab_size=cfg.type_vocab_size, layer_norm_eps=1e-12,
        pad_token_id=0, hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1, **extra)
    c._attn_implementation = "eager"
    return c


@pytest.mark.parametrize("d,ffn,layers", [(96, 192, 2), (768, 3072, 2)])
def test_shim_encoder_equals_installed_transformers_bert(d, ffn, layers):
    torch.manual_seed(0)
    cfg = R.bert_config(300, 2, layers, d=d, ffn=ffn, heads=12)
    enc = R.EncoderModel(cfg).eval()
    hf_model = transformers.BertModel(_hf_config(cfg), add_pooling_layer=False).eval()
    res = hf_model.load_state_dict(enc.state_dict(), strict=False)
    assert not res.missing_keys and not res.unexpected_keys         # identical state-dict key names (Appendix A.10)
    ids = torch.randint(5, 300, (3, 11))
    mask = torch.ones(3, 11, dtype=torch.long)
    mask[1, 7:] = 0
    mask[2, 4:] = 0
    types = torch.tensor([[0] * 11, [1] * 11, [0] * 11])
    with torch.no_grad():
        a = enc(ids, attention_mask=mask, token_type_ids=types)
        b = hf_model(input_ids=ids, attention_mask=mask, token_type_ids=types).last_hidden_state
    err = float((a - b)[mask.bool()].abs().max())
    assert err < (2e-6 if d == 96 else 1e-5), err


@pytest.mark.parametrize("d,ffn,layers", [(96, 192, 2), (768, 3072, 2)])
def test_shim_causal_cross_attention_decoder_equals_installed_transformers_bert(d, ffn, layers):
    torch.manual_seed(1)
    cfg = R.bert_config(300, 2, layers, d=d, ffn=ffn, heads=12)
    dec = R.DecoderModel(cfg).eval()
    hf_model = transformers.BertModel(_hf_config(cfg, is_decoder=True, add_cross_attention=True), add_pooling_layer=False).eval()
    sd = {k.replace("decoder.layer", "encoder.layer"): v for k, v in dec.state_dict().items()}
    res = hf_model.load_state_dict(sd, strict=False)
    assert not res.missing_keys and not res.unexpected_keys
    mem = torch.randn(3, 9, d)
    mmask = torch.ones(3, 9)
    mmask[0, 5:] = 0
    tgt = torch.randint(5, 300, (3, 6))
    types = torch.ones_like(tgt)
    with torch.no_grad():
        a = dec(input_ids=tgt, encoder_attention_mask=mmask, token_type_ids=types, encoder_states=mem)
        b = hf_model(input_ids=tgt, token_type_ids=types, encoder_hidden_states=mem, encoder_attention_mask=mmask,
                     use_cache=False).last_hidden_state
    err = float((a - b).abs().max())
    assert err < (2e-6 if d == 96 else 1e-5), err


def test_causal_prefix_invariance():
    """The decoder state at position p depends on tokens 0..p only (causal mask), so the last-position state of a
    prefix equals that position's state inside any longer sequence: what makes decoding with a KV cache exact."""
    torch.manual_seed(2)
    cfg = R.bert_config(400, 2, 3, d=768, ffn=3072, heads=12)
    dec = R.DecoderModel(cfg).eval()
    mem = torch.randn(2, 7, 768)
    mmask = torch.ones(2, 7)
    mmask[1, 4:] = 0
    long = torch.randint(5, 400, (2, 9))
    with torch.no_grad():
        full = dec(input_ids=long, encoder_attention_mask=mmask, token_type_ids=torch.ones_like(long), encoder_states=mem)
        for n in (1, 5, 8):
            part = dec(input_ids=long[:, :n], encoder_attention_mask=mmask, token_type_ids=torch.ones_like(long[:, :n]),
                       encoder_states=mem)
            assert float((part - full[:, :n]).abs().max()) < 5e-6
