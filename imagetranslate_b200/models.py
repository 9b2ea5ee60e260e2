"""Model containers for the decode path: parameter-name compatible with the reference checkpoints
(reference src/seq2seq.py:183-213 `mt_model.state_dict`, key list in SURVEY.md Appendix A.10), so a
released ImageTranslate state dict loads with `load_state_dict`.

Only what decoding needs is implemented: parameter storage for the decoder(s) / output layer(s) (consumed
by engine.PackedDecoder), and the bidirectional encoder forward that produces the memory
(reference src/seq2seq.py:101-108 -> src/bert_seq2seq.py:103-144).  Training forwards are out of scope.
"""
import copy
import math
import os
import pickle

import torch
import torch.nn as nn
import torch.nn.functional as F


class BertConfig(object):
    def __init__(self, vocab_size, type_vocab_size, num_hidden_layers=6, hidden_size=768, intermediate_size=3072,
                 num_attention_heads=12, max_position_embeddings=512, pad_token_id=0, bos_token_id=1, eos_token_id=4,
                 layer_norm_eps=1e-12, initializer_range=0.02, is_decoder=False):
        self.__dict__.update({k: v for k, v in locals().items() if k != "self"})


class Embeddings(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.word_embeddings = nn.Embedding(c.vocab_size, c.hidden_size, padding_idx=c.pad_token_id)
        self.position_embeddings = nn.Embedding(c.max_position_embeddings, c.hidden_size)
        self.token_type_embeddings = nn.Embedding(c.type_vocab_size, c.hidden_size)
        self.LayerNorm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)

    def forward(self, ids, type_ids):
        pos = torch.arange(ids.size(1), device=ids.device)
        return self.LayerNorm(self.word_embeddings(ids) + self.position_embeddings(pos)[None] + self.token_type_embeddings(type_ids))


class SelfAttention(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.num_attention_heads = c.num_attention_heads
        self.query = nn.Linear(c.hidden_size, c.hidden_size)
        self.key = nn.Linear(c.hidden_size, c.hidden_size)
        self.value = nn.Linear(c.hidden_size, c.hidden_size)

    def forward(self, x, add_mask):
        B, S, d = x.shape
        H = self.num_attention_heads
        sp = lambda t: t.view(B, S, H, d // H).transpose(1, 2)
        o = F.scaled_dot_product_attention(sp(self.query(x)), sp(self.key(x)), sp(self.value(x)), attn_mask=add_mask)
        return o.transpose(1, 2).reshape(B, S, d)


class DenseLN(nn.Module):
    def __init__(self, n_in, n_out, eps):
        super().__init__()
        self.dense = nn.Linear(n_in, n_out)
        self.LayerNorm = nn.LayerNorm(n_out, eps=eps)

    def forward(self, h, residual):
        return self.LayerNorm(self.dense(h) + residual)


class Attention(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.self = SelfAttention(c)
        self.output = DenseLN(c.hidden_size, c.hidden_size, c.layer_norm_eps)

    def forward(self, x, add_mask):
        return self.output(self.self(x, add_mask), x)


class Intermediate(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.dense = nn.Linear(c.hidden_size, c.intermediate_size)

    def forward(self, x):
        return F.gelu(self.dense(x))


class Layer(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.attention = Attention(c)
        if c.is_decoder:
            self.crossattention = Attention(c)
        self.intermediate = Intermediate(c)
        self.output = DenseLN(c.intermediate_size, c.hidden_size, c.layer_norm_eps)

    def forward(self, x, add_mask):
        a = self.attention(x, add_mask)
        return self.output(self.intermediate(a), a)


class Stack(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.layer = nn.ModuleList([Layer(c) for _ in range(c.num_hidden_layers)])


def _init_bert(module, std):
    """HF 2.9.0 init_weights."""
    for m in module.modules():
        if isinstance(m, (nn.Linear, nn.Embedding)):
            m.weight.data.normal_(0.0, std)
        if isinstance(m, nn.LayerNorm):
            m.weight.data.fill_(1.0)
            m.bias.data.zero_()
        if isinstance(m, nn.Linear) and m.bias is not None:
            m.bias.data.zero_()


class BertEncoderModel(nn.Module):
    """Bidirectional encoder (reference src/bert_seq2seq.py:94-144)."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.embeddings = Embeddings(config)
        self.encoder = Stack(config)
        _init_bert(self, config.initializer_range)

    def forward(self, input_ids, attention_mask=None, token_type_ids=None):
        x = self.embeddings(input_ids, token_type_ids)
        add_mask = None
        if attention_mask is not None:
            add_mask = (1.0 - attention_mask[:, None, None, :].to(x.dtype)) * -10000.0
        for ly in self.encoder.layer:
            x = ly(x, add_mask)
        return x


class BertDecoderModel(nn.Module):
    """Parameter container of the causal decoder with cross-attention (reference src/bert_seq2seq.py:15-91).
    Its forward is the CUDA decode path (imagetranslate_b200.BeamDecoder)."""

    def __init__(self, config):
        super().__init__()
        self.config = copy.deepcopy(config)
        self.config.is_decoder = True
        self.embeddings = Embeddings(self.config)
        self.decoder = Stack(self.config)
        _init_bert(self, config.initializer_range)


class BertOutputLayer(nn.Module):
    """reference src/bert_seq2seq.py:6-12"""

    def __init__(self, config):
        super().__init__()
        self.layer = nn.Linear(config.hidden_size, config.vocab_size)

    def forward(self, x):
        return self.layer(x)


class Seq2Seq(nn.Module):
    """reference src/seq2seq.py:20-86 (structure + sharing), :101-108 (encode), :183-213 (save/load)."""

    def __init__(self, text_processor, lang_dec: bool = True, use_proposals=False, tie_embed=False, enc_layer: int = 6,
                 dec_layer: int = 3, embed_dim: int = 768, intermediate_dim: int = 3072, freeze_image: bool = False,
                 resnet_depth: int = 1, use_obj: bool = False):
        super().__init__()
        if use_proposals:
            raise RuntimeError("use_proposals is not supported by the B200 decode path yet")
        self.text_processor = text_processor
        langs = text_processor.languages
        self.config = BertConfig(vocab_size=text_processor.tokenizer.get_vocab_size(), type_vocab_size=len(langs),
                                 num_hidden_layers=enc_layer, hidden_size=embed_dim, intermediate_size=intermediate_dim,
                                 pad_token_id=text_processor.pad_token_id(), bos_token_id=text_processor.bos_token_id(),
                                 eos_token_id=text_processor.sep_token_id())
        dec_config = copy.deepcopy(self.config)
        dec_config.num_hidden_layers = dec_layer
        self.enc_layer, self.dec_layer, self.embed_dim, self.intermediate_dim = enc_layer, dec_layer, embed_dim, intermediate_dim
        self.lang_dec, self.tie_embed, self.use_proposals = lang_dec, tie_embed, use_proposals
        self.freeze_image, self.resnet_depth = freeze_image, resnet_depth
        self.encoder = BertEncoderModel(self.config)
        ee = self.encoder.embeddings
        if not lang_dec:
            self.decoder = BertDecoderModel(dec_config)
            de = self.decoder.embeddings
            de.position_embeddings.weight = ee.position_embeddings.weight
            de.token_type_embeddings.weight = ee.token_type_embeddings.weight
            de.word_embeddings.weight = ee.word_embeddings.weight
            if tie_embed:
                self.output_layer = BertOutputLayer(dec_config)
                # :56/:59 register the embedding as `output_layer.weight` on the WRAPPER; the projection
                # `output_layer.layer.weight` stays untied.  Kept so that state-dict keys match.
                self.output_layer.weight = de.word_embeddings.weight
            else:
                self.output_layer = nn.ModuleList([BertOutputLayer(dec_config) for _ in langs])
            if enc_layer == dec_layer:                               # :63-65 shared self-attention blocks
                for i in range(enc_layer):
                    self.decoder.decoder.layer[i].attention = self.encoder.encoder.layer[i].attention
        else:
            dec = BertDecoderModel(dec_config)
            self.decoder = nn.ModuleList([copy.deepcopy(dec) for _ in langs])
            self.output_layer = nn.ModuleList([BertOutputLayer(dec_config) for _ in langs])
            for i, dec in enumerate(self.decoder):
                if tie_embed:
                    dec.embeddings.position_embeddings = ee.position_embeddings
                self.output_layer[i].weight = dec.embeddings.word_embeddings.weight     # :75 (wrapper alias, see above)
                # :76-77 _tie_or_clone_weights(out=encoder..., inp=dec...) sets out.weight = inp.weight: the ENCODER's
                # table is re-pointed at each decoder's in turn, so it ends up shared with the LAST decoder only and
                # every other decoder keeps its own (deep-copied) token-type table.
                ee.token_type_embeddings.weight = dec.embeddings.token_type_embeddings.weight

    def encode(self, src_inputs, src_mask, src_langs, images=None):
        dev = self.encoder.embeddings.word_embeddings.weight.device
        return (self.encoder(src_inputs.to(dev), attention_mask=src_mask.to(dev), token_type_ids=src_langs.to(dev)), None)

    def save(self, out_dir):
        os.makedirs(out_dir, exist_ok=True)
        with open(os.path.join(out_dir, "mt_config"), "wb") as fp:
            pickle.dump((self.lang_dec, self.use_proposals, self.enc_layer, self.dec_layer, self.embed_dim,
                         self.intermediate_dim, self.tie_embed, self.resnet_depth, self.freeze_image), fp)
        torch.save(self.state_dict(), os.path.join(out_dir, "mt_model.state_dict"))

    @staticmethod
    def load(cls, out_dir, text_processor, map_location="cpu"):
        with open(os.path.join(out_dir, "mt_config"), "rb") as fp:
            lang_dec, use_proposals, enc_layer, dec_layer, embed_dim, intermediate_dim, tie_embed, resnet_depth, freeze_image = pickle.load(fp)
        m = cls(text_processor, lang_dec=lang_dec, use_proposals=use_proposals, tie_embed=tie_embed, enc_layer=enc_layer,
                dec_layer=dec_layer, embed_dim=embed_dim, intermediate_dim=intermediate_dim, freeze_image=freeze_image,
                resnet_depth=resnet_depth)
        m.load_state_dict(torch.load(os.path.join(out_dir, "mt_model.state_dict"), map_location=map_location), strict=False)
        return m


class MassSeq2Seq(Seq2Seq):
    """reference src/mass_seq2seq.py -- identical decode surface."""


class SyntheticTextProcessor(object):
    """Minimal stand-in for reference src/textprocessor.py when no tokenizer files exist (benchmarks):
    pad 0, bos 1, unk 2, mask 3, </s> 4, language tokens from 5 (src/textprocessor.py:24-30)."""

    class _Tok(object):
        def __init__(self, v):
            self._v = v

        def get_vocab_size(self):
            return self._v

    def __init__(self, vocab_size, languages=("<en>", "<fa>")):
        self.tokenizer = SyntheticTextProcessor._Tok(vocab_size)
        self.languages = {l: i for i, l in enumerate(languages)}

    def pad_token_id(self):
        return 0

    def bos_token_id(self):
        return 1

    def sep_token_id(self):
        return 4

    def token_id(self, tok):
        return 5 + self.languages[tok]
