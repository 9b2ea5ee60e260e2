"""ORACLE / TEST INFRASTRUCTURE ONLY -- restatement of the BERT arithmetic of
`transformers==2.9.0` (third-party, pinned at /root/reference/src/requirements.txt:8,
NOT vendored under /root/reference).  Parity anchor: state-dict keys and outputs
match the installed transformers-5.x eager BertModel (tests/test_oracle_anchor.py).

Reference call sites this serves: src/bert_seq2seq.py:15-23,69-71,79-90,94-101,130-143;
src/seq2seq.py:37,41-59,73-77.  The module namespace also has to export torch/nn/F/
math/copy because the reference star-imports it (src/bert_seq2seq.py:3).
"""
import copy  # noqa: F401
import math

import torch
import torch.nn as nn
import torch.nn.functional as F  # noqa: F401


class BertConfig(object):
    """Attribute bag with the 2.9.0 defaults the reference relies on (Appendix A.1)."""

    _defaults = dict(
        vocab_size=30522, hidden_size=768, num_hidden_layers=12, num_attention_heads=12,
        intermediate_size=3072, hidden_act="gelu", hidden_dropout_prob=0.1,
        attention_probs_dropout_prob=0.1, max_position_embeddings=512, type_vocab_size=2,
        initializer_range=0.02, layer_norm_eps=1e-12, pad_token_id=0, bos_token_id=None,
        eos_token_id=None, is_decoder=False, output_attentions=False, output_hidden_states=False,
    )

    def __init__(self, **kwargs):
        for k, v in self._defaults.items():
            setattr(self, k, v)
        for k, v in kwargs.items():
            setattr(self, k, v)

    def to_dict(self):
        return dict(self.__dict__)

    def __eq__(self, other):
        return isinstance(other, BertConfig) and self.__dict__ == other.__dict__

    def __repr__(self):
        return "BertConfig(%r)" % (self.__dict__,)


def _gelu_erf(x):
    return x * 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0)))


class BertPreTrainedModel(nn.Module):
    """Helpers of PreTrainedModel / ModuleUtilsMixin used by the reference (Appendix A.8)."""

    def __init__(self, config, *a, **kw):
        super().__init__()
        self.config = config

    @property
    def device(self):
        return next(self.parameters()).device

    @property
    def dtype(self):
        return next(self.parameters()).dtype

    def _init_weights(self, module):
        if isinstance(module, (nn.Linear, nn.Embedding)):
            module.weight.data.normal_(mean=0.0, std=self.config.initializer_range)
        elif isinstance(module, nn.LayerNorm):
            module.bias.data.zero_()
            module.weight.data.fill_(1.0)
        if isinstance(module, nn.Linear) and module.bias is not None:
            module.bias.data.zero_()

    def init_weights(self):
        self.apply(self._init_weights)

    def _tie_or_clone_weights(self, output_embeddings, input_embeddings):
        output_embeddings.weight = input_embeddings.weight
        if getattr(output_embeddings, "bias", None) is not None:
            output_embeddings.bias.data = torch.nn.functional.pad(
                output_embeddings.bias.data,
                (0, output_embeddings.weight.shape[0] - output_embeddings.bias.shape[0]),
                "constant", 0)
        if hasattr(output_embeddings, "out_features") and hasattr(input_embeddings, "num_embeddings"):
            output_embeddings.out_features = input_embeddings.num_embeddings

    def get_head_mask(self, head_mask, num_hidden_layers, is_attention_chunked=False):
        assert head_mask is None
        return [None] * num_hidden_layers

    def get_extended_attention_mask(self, attention_mask, input_shape, device):
        if attention_mask.dim() == 3:
            ext = attention_mask[:, None, :, :]
        elif attention_mask.dim() == 2:
            if self.config.is_decoder:
                batch_size, seq_length = input_shape
                seq_ids = torch.arange(seq_length, device=device)
                causal = seq_ids[None, None, :].repeat(batch_size, seq_length, 1) <= seq_ids[None, :, None]
                causal = causal.to(attention_mask.dtype)
                ext = causal[:, None, :, :] * attention_mask[:, None, None, :]
            else:
                ext = attention_mask[:, None, None, :]
        else:
            raise ValueError("bad attention_mask shape")
        ext = ext.to(dtype=self.dtype)
        return (1.0 - ext) * -10000.0


class BertEmbeddings(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.word_embeddings = nn.Embedding(config.vocab_size, config.hidden_size, padding_idx=config.pad_token_id)
        self.position_embeddings = nn.Embedding(config.max_position_embeddings, config.hidden_size)
        self.token_type_embeddings = nn.Embedding(config.type_vocab_size, config.hidden_size)
        self.LayerNorm = nn.LayerNorm(config.hidden_size, eps=config.layer_norm_eps)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)

    def forward(self, input_ids=None, token_type_ids=None, position_ids=None, inputs_embeds=None):
        shape = input_ids.size() if input_ids is not None else inputs_embeds.size()[:-1]
        seq_length = shape[1]
        device = input_ids.device if input_ids is not None else inputs_embeds.device
        if position_ids is None:
            position_ids = torch.arange(seq_length, dtype=torch.long, device=device).unsqueeze(0).expand(shape)
        if token_type_ids is None:
            token_type_ids = torch.zeros(shape, dtype=torch.long, device=device)
        if inputs_embeds is None:
            inputs_embeds = self.word_embeddings(input_ids)
        e = inputs_embeds + self.position_embeddings(position_ids) + self.token_type_embeddings(token_type_ids)
        return self.dropout(self.LayerNorm(e))


class BertSelfAttention(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.num_attention_heads = config.num_attention_heads
        self.attention_head_size = config.hidden_size // config.num_attention_heads
        self.all_head_size = self.num_attention_heads * self.attention_head_size
        self.query = nn.Linear(config.hidden_size, self.all_head_size)
        self.key = nn.Linear(config.hidden_size, self.all_head_size)
        self.value = nn.Linear(config.hidden_size, self.all_head_size)
        self.dropout = nn.Dropout(config.attention_probs_dropout_prob)

    def _split(self, x):
        return x.view(x.size(0), x.size(1), self.num_attention_heads, self.attention_head_size).permute(0, 2, 1, 3)

    def forward(self, hidden_states, attention_mask=None, head_mask=None, encoder_hidden_states=None,
                encoder_attention_mask=None):
        q = self.query(hidden_states)
        if encoder_hidden_states is not None:
            k = self.key(encoder_hidden_states)
            v = self.value(encoder_hidden_states)
            attention_mask = encoder_attention_mask
        else:
            k = self.key(hidden_states)
            v = self.value(hidden_states)
        q, k, v = self._split(q), self._split(k), self._split(v)
        scores = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(self.attention_head_size)
        if attention_mask is not None:
            scores = scores + attention_mask
        probs = self.dropout(nn.Softmax(dim=-1)(scores))
        ctx = torch.matmul(probs, v).permute(0, 2, 1, 3).contiguous()
        return (ctx.view(ctx.size(0), ctx.size(1), self.all_head_size),)


class BertSelfOutput(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dense = nn.Linear(config.hidden_size, config.hidden_size)
        self.LayerNorm = nn.LayerNorm(config.hidden_size, eps=config.layer_norm_eps)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)

    def forward(self, hidden_states, input_tensor):
        return self.LayerNorm(self.dropout(self.dense(hidden_states)) + input_tensor)


class BertAttention(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.self = BertSelfAttention(config)
        self.output = BertSelfOutput(config)

    def forward(self, hidden_states, attention_mask=None, head_mask=None, encoder_hidden_states=None,
                encoder_attention_mask=None):
        s = self.self(hidden_states, attention_mask, head_mask, encoder_hidden_states, encoder_attention_mask)
        return (self.output(s[0], hidden_states),)


class BertIntermediate(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dense = nn.Linear(config.hidden_size, config.intermediate_size)
        assert config.hidden_act == "gelu"

    def forward(self, hidden_states):
        return _gelu_erf(self.dense(hidden_states))


class BertOutput(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dense = nn.Linear(config.intermediate_size, config.hidden_size)
        self.LayerNorm = nn.LayerNorm(config.hidden_size, eps=config.layer_norm_eps)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)

    def forward(self, hidden_states, input_tensor):
        return self.LayerNorm(self.dropout(self.dense(hidden_states)) + input_tensor)


class BertLayer(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.attention = BertAttention(config)
        self.is_decoder = config.is_decoder
        if self.is_decoder:
            self.crossattention = BertAttention(config)
        self.intermediate = BertIntermediate(config)
        self.output = BertOutput(config)

    def forward(self, hidden_states, attention_mask=None, head_mask=None, encoder_hidden_states=None,
                encoder_attention_mask=None):
        a = self.attention(hidden_states, attention_mask, head_mask)[0]
        if self.is_decoder and encoder_hidden_states is not None:
            a = self.crossattention(a, attention_mask, head_mask, encoder_hidden_states, encoder_attention_mask)[0]
        return (self.output(self.intermediate(a), a),)


class BertEncoder(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.layer = nn.ModuleList([BertLayer(config) for _ in range(config.num_hidden_layers)])

    def forward(self, hidden_states, attention_mask=None, head_mask=None, encoder_hidden_states=None,
                encoder_attention_mask=None):
        for i, layer_module in enumerate(self.layer):
            hidden_states = layer_module(hidden_states, attention_mask, None, encoder_hidden_states,
                                         encoder_attention_mask)[0]
        return (hidden_states,)


class BertModel(BertPreTrainedModel):
    """Only referenced for type annotations in src/lm.py:15-16."""

    def __init__(self, config):
        super().__init__(config)
        self.embeddings = BertEmbeddings(config)
        self.encoder = BertEncoder(config)
        self.init_weights()
