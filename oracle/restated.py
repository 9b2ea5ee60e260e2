"""ORACLE -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Plain-PyTorch fp32 restatement of the reference decode path.  It can travel to
the GPU box (where /root/reference does not exist).  It is validated here against
the real reference modules (tests/test_oracle_vs_reference.py) and against the
committed golden fixtures (tests/test_oracle_golden.py).

Restated pieces and the reference lines they follow:
  * FakeTextProcessor          -- attribute surface of src/textprocessor.py:22-62 used at
                                   src/seq2seq.py:26-36 and src/seq_gen.py:107
  * OutputLayer / DecoderModel / EncoderModel
                               -- src/bert_seq2seq.py:6-12, 15-91, 94-144
  * Seq2SeqModel               -- src/seq2seq.py:20-86 (construction, sharing), :101-108 (encode);
                                   src/mass_seq2seq.py (same decode surface)
  * beam_search                -- src/seq_gen.py:46-242 (BeamDecoder.forward), :6-24, :37-44
Like the reference this recomputes the whole prefix every step (no KV cache).
"""
import copy
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .shim.transformers import modeling_bert as hf


class FakeTextProcessor(object):
    """pad 0, bos 1, unk 2, mask 3, sep 4, languages from 5 (src/textprocessor.py:24-30)."""

    class _Tok(object):
        def __init__(self, v):
            self._v = v

        def get_vocab_size(self):
            return self._v

    def __init__(self, vocab_size, languages=("<en>", "<fa>")):
        self.tokenizer = FakeTextProcessor._Tok(vocab_size)
        self.languages = {l: i for i, l in enumerate(languages)}
        self._lang_tok = {l: 5 + i for i, l in enumerate(languages)}

    def pad_token_id(self):
        return 0

    def bos_token_id(self):
        return 1

    def sep_token_id(self):
        return 4

    def token_id(self, tok):
        return self._lang_tok[tok]

    def vocab_size(self):
        return self.tokenizer.get_vocab_size()


def bert_config(vocab_size, n_lang, layers, d=768, ffn=3072, heads=12, pad=0, bos=1, eos=4):
    """src/lm_config.py:4-30 + src/seq2seq.py:36."""
    return hf.BertConfig(
        attention_probs_dropout_prob=0.1, hidden_act="gelu", hidden_dropout_prob=0.1, hidden_size=d,
        initializer_range=0.02, intermediate_size=ffn, max_position_embeddings=512,
        num_attention_heads=heads, num_hidden_layers=layers, vocab_size=vocab_size, pad_token_id=pad,
        bos_token_id=bos, eos_token_id=eos, type_vocab_size=n_lang)


class OutputLayer(nn.Module):
    """src/bert_seq2seq.py:6-12 -- nn.Linear(d, V) with bias, default torch init."""

    def __init__(self, config):
        super().__init__()
        self.layer = nn.Linear(config.hidden_size, config.vocab_size)

    def forward(self, x):
        return self.layer(x)


class EncoderModel(hf.BertPreTrainedModel):
    """src/bert_seq2seq.py:94-144."""

    def __init__(self, config):
        super().__init__(config)
        self.embeddings = hf.BertEmbeddings(config)
        self.encoder = hf.BertEncoder(config)
        self.init_weights()

    def forward(self, input_ids, attention_mask=None, token_type_ids=None):
        if attention_mask is None:
            attention_mask = torch.ones(input_ids.size(), device=input_ids.device)
        ext = self.get_extended_attention_mask(attention_mask, input_ids.size(), self.device)
        emb = self.embeddings(input_ids=input_ids, token_type_ids=token_type_ids)
        return self.encoder(emb, attention_mask=ext)[0]


class DecoderModel(hf.BertPreTrainedModel):
    """src/bert_seq2seq.py:15-91 (is_decoder=True: causal mask + cross-attention)."""

    def __init__(self, config):
        super().__init__(config)
        self.config = copy.deepcopy(config)
        self.config.is_decoder = True
        self.embeddings = hf.BertEmbeddings(self.config)
        self.decoder = hf.BertEncoder(self.config)
        self.init_weights()

    def forward(self, input_ids=None, encoder_attention_mask=None, tgt_attention_mask=None, token_type_ids=None,
                encoder_states=None):
        dev = input_ids.device
        if tgt_attention_mask is None:
            tgt_attention_mask = torch.ones(input_ids.size(), device=dev)
        ext = self.get_extended_attention_mask(tgt_attention_mask, input_ids.size(), self.device)
        if encoder_attention_mask is None:
            encoder_attention_mask = torch.ones(encoder_states.size()[:2], device=dev)
        # src/bert_seq2seq.py:25-38
        enc_ext = (1.0 - encoder_attention_mask[:, None, None, :].to(dtype=self.dtype)) * -10000.0
        emb = self.embeddings(input_ids=input_ids, token_type_ids=token_type_ids)
        return self.decoder(emb, attention_mask=ext, encoder_hidden_states=encoder_states,
                            encoder_attention_mask=enc_ext)[0]


class Seq2SeqModel(nn.Module):
    """Decode-relevant restatement of src/seq2seq.py:20-86,101-108.  Parameter names equal the
    reference's state-dict keys, so state dicts interchange."""

    def __init__(self, text_processor, lang_dec=True, use_proposals=False, tie_embed=False, enc_layer=6,
                 dec_layer=3, embed_dim=768, intermediate_dim=3072):
        super().__init__()
        self.text_processor = text_processor
        n_lang = len(text_processor.languages)
        self.config = bert_config(text_processor.tokenizer.get_vocab_size(), n_lang, enc_layer, embed_dim,
                                  intermediate_dim, pad=text_processor.pad_token_id(),
                                  bos=text_processor.bos_token_id(), eos=text_processor.sep_token_id())
        dec_config = copy.deepcopy(self.config)
        dec_config.num_hidden_layers = dec_layer
        self.enc_layer, self.dec_layer = enc_layer, dec_layer
        self.embed_dim, self.intermediate_dim = embed_dim, intermediate_dim
        self.encoder = EncoderModel(self.config)
        self.encoder.init_weights()
        self.lang_dec, self.tie_embed, self.use_proposals = lang_dec, tie_embed, use_proposals
        tie = self.encoder._tie_or_clone_weights
        if not lang_dec:  # src/seq2seq.py:45-65
            self.decoder = DecoderModel(dec_config)
            tie(self.encoder.embeddings.position_embeddings, self.decoder.embeddings.position_embeddings)
            tie(self.encoder.embeddings.token_type_embeddings, self.decoder.embeddings.token_type_embeddings)
            tie(self.encoder.embeddings.word_embeddings, self.decoder.embeddings.word_embeddings)
            if tie_embed:
                self.output_layer = OutputLayer(dec_config)
                # :56/:59 set `.weight` on the WRAPPER, not on `.layer` -- the projection stays untied.
                tie(self.output_layer, self.encoder.embeddings.word_embeddings)
                tie(self.encoder.embeddings.position_embeddings, self.decoder.embeddings.position_embeddings)
                self.decoder._tie_or_clone_weights(self.output_layer, self.decoder.embeddings.word_embeddings)
            else:
                self.output_layer = nn.ModuleList([OutputLayer(dec_config) for _ in text_processor.languages])
            if len(self.encoder.encoder.layer) == len(self.decoder.decoder.layer):
                for i in range(len(self.encoder.encoder.layer)):
                    self.decoder.decoder.layer[i].attention = self.encoder.encoder.layer[i].attention
        else:  # src/seq2seq.py:67-77
            dec = DecoderModel(dec_config)
            self.decoder = nn.ModuleList([copy.deepcopy(dec) for _ in text_processor.languages])
            self.output_layer = nn.ModuleList([OutputLayer(dec_config) for _ in text_processor.languages])
            for i, dec in enumerate(self.decoder):
                if tie_embed:
                    tie(self.output_layer[i], self.encoder.embeddings.word_embeddings)
                    dec.embeddings.position_embeddings = self.encoder.embeddings.position_embeddings
                dec._tie_or_clone_weights(self.output_layer[i], dec.embeddings.word_embeddings)
                dec._tie_or_clone_weights(self.encoder.embeddings.token_type_embeddings,
                                          dec.embeddings.token_type_embeddings)

        if use_proposals:  # src/seq2seq.py:79-83
            self.proposal_embedding = self.encoder.embeddings.word_embeddings
            self.lexical_gate = nn.Parameter(torch.zeros(1, self.config.hidden_size).fill_(0.1), requires_grad=True)
            self.lexical_layer_norm = nn.LayerNorm(self.config.hidden_size, eps=self.config.layer_norm_eps)

    def encode(self, src_inputs, src_mask, src_langs, images=None):
        dev = self.encoder.embeddings.word_embeddings.weight.device
        return (self.encoder(src_inputs.to(dev), attention_mask=src_mask.to(dev),
                             token_type_ids=src_langs.to(dev)), None)

    def attend_proposal(self, decoder_output, proposals, pad_idx):
        """src/seq2seq.py:110-144 for the 2-D (last state) input of the decode path (:126-131 branch).  Line 136,
        `attend_scores[attend_mask].fill_(-10000.0)`, fills a COPY (boolean-mask indexing): PAD proposals stay attended."""
        dev = self.encoder.embeddings.word_embeddings.weight.device
        proposals = proposals.to(dev)
        emb = self.proposal_embedding(proposals)                              # (B,P,d)
        if len(proposals) < len(decoder_output):
            beam = int(len(decoder_output) / len(proposals))
            proposals = torch.repeat_interleave(proposals, beam, 0)
            emb = torch.repeat_interleave(emb, beam, 0)
        scores = torch.matmul(decoder_output.unsqueeze(1), emb.transpose(-1, -2)).squeeze(-2)   # (M,P), unscaled, unmasked
        probs = torch.softmax(scores, dim=-1)
        ctx = torch.sum(probs.unsqueeze(-1) * emb, dim=-2)
        ctx[torch.all(proposals == pad_idx, dim=-1)] = 1e-8
        sig = torch.sigmoid(self.lexical_gate + 1e-8)
        return self.lexical_layer_norm(sig * decoder_output + (1 - sig) * ctx)


def len_penalty(lengths, alpha):
    """src/seq_gen.py:37-44."""
    return torch.pow((lengths + 6.0) / 6.0, alpha).unsqueeze(-1)


def outputs_until_eos(eos, outputs, size_limit=None):
    """src/seq_gen.py:6-24: cut each row before its first EOS, else at size_limit[r]."""
    outputs = outputs.cpu()
    if outputs.dim() == 1:
        outputs = outputs.unsqueeze(0)
    res = []
    for r in range(outputs.size(0)):
        hits = (outputs[r] == eos).nonzero()
        if hits.numel() > 0:
            res.append(outputs[r, :int(hits[0, 0])])
        else:
            last = outputs.size(1) if size_limit is None else int(size_limit[r])
            res.append(outputs[r, :last])
    return res


def search_step(i, log_probs, prefixes, scores, sizes, max_lens, eos, pad_idx, beam, alpha, topk_fn=None):
    """One iteration of src/seq_gen.py:131-227 AFTER the decoder call, i.e. lines 138-143 and 194-227.

    log_probs : (M, V) f32 log-softmax rows (M = B at i == 1 else B*beam)
    prefixes  : (B, 1 or beam, i) int64;  scores (B, 1 or beam) f32;  sizes (B,) / (B, beam) f32 or None
    Returns new (prefixes, scores, sizes, raw_topk_indices).
    """
    B = prefixes.size(0)
    V = log_probs.size(-1)
    cur = prefixes.view(-1, prefixes.size(-1))
    fin = torch.any(cur == eos, 1)                                  # :143
    lim = (max_lens < (i + 1)).unsqueeze(-1).expand(-1, beam)        # :138-139
    out = log_probs.clone()
    out[fin] = 0                                                     # :195
    if i > 1:
        out[lim.contiguous().view(-1)] = 0                           # :197
    cur_scores = scores.view(-1).unsqueeze(-1)
    if beam > 1:
        cand = ((cur_scores + out) / len_penalty(sizes.view(-1), alpha)).view(B, -1)   # :199
    else:
        cand = (cur_scores + out).view(B, -1)                        # :201
    if topk_fn is None:
        top, idx = torch.topk(cand, k=beam, dim=1)                   # :203
    else:
        top, idx = topk_fn(cand, beam)
    raw_idx = idx.clone()
    if i > 1:
        idx[lim] = pad_idx                                           # :207
        flat = idx.view(-1)
        flat[fin] = pad_idx                                          # :213 (indexed by NEW slot, OLD flag)
        src = idx // V                                               # :216 (integer division, torch<=1.4)
        pre = prefixes.gather(1, src.unsqueeze(-1).expand(-1, -1, prefixes.size(-1))).view(-1, i)   # :217-218
        sz = sizes.gather(1, src).view(-1) if beam > 1 else None     # :219
    else:
        pre = torch.repeat_interleave(prefixes, beam, 0).view(-1, i)  # :221
        sz = torch.repeat_interleave(sizes, beam, 0) if beam > 1 else None
    words = (idx.view(-1) % V).unsqueeze(-1)                         # :223 (vocab[flat_indices])
    new_prefixes = torch.cat([pre, words], dim=1).view(B, beam, i + 1)   # :224
    new_sizes = None
    if beam > 1:
        new_sizes = (sz + (~(words.squeeze(-1) == pad_idx)).float()).view(B, beam)     # :226
    return new_prefixes, top, new_sizes, raw_idx


def lowest_index_topk(cand, k):
    """Deterministic tie order used by the CUDA path: descending value, then ascending flat index."""
    B, N = cand.shape
    vals, idxs = [], []
    for b in range(B):
        order = sorted(range(N), key=lambda j: (-float(cand[b, j]), j))[:k] if N <= 4096 else None
        if order is None:
            v, _ = torch.topk(cand[b], k)
            thr = v[-1]
            pool = (cand[b] >= thr).nonzero().view(-1).tolist()
            order = sorted(pool, key=lambda j: (-float(cand[b, j]), j))[:k]
        idxs.append(order)
        vals.append([float(cand[b, j]) for j in order])
    return torch.tensor(vals, dtype=cand.dtype), torch.tensor(idxs, dtype=torch.long)


@torch.no_grad()
def beam_search(model, src_inputs=None, src_sizes=None, first_tokens=None, src_mask=None, src_langs=None,
                tgt_langs=None, pad_idx=None, max_len=None, unpad_output=True, beam_width=5, image_embed=None,
                max_len_a=1.1, max_len_b=5, len_penalty_ratio=0.8, trace=None, topk_fn=None,
                teacher_prefixes=None, proposals=None, memory2=None, decoder2=None, gate=None):
    """src/seq_gen.py:46-242 for the text (`src_inputs`) and `image_embed` memories.

    Second decoder stream (:167-188): `memory2` (B,S2,d) is decoded by `decoder2` (obj_decoder; or the model's own decoder
    for the text+image mode) with no memory mask, and mixed by sigmoid(gate + 1e-7).  `proposals` (B,P): :191-192.

    trace: optional list; per step a dict(i, last_state, log_probs, prefixes, scores, sizes, raw_idx)
    is appended (the recorder of SURVEY Appendix D.6).
    """
    device = model.encoder.embeddings.word_embeddings.weight.device
    beam = beam_width
    batch_lang = int(tgt_langs[0])
    if src_inputs is not None:
        B = src_inputs.size(0)
        src_mask = src_mask.to(device)
        langs2d = src_langs.unsqueeze(-1).expand(-1, src_inputs.size(-1))
        memory = model.encode(src_inputs, src_mask, langs2d)[0]          # :95-96
    else:
        B = image_embed.size(0)
        memory = image_embed.to(device)                                   # :100-103
    eos = model.text_processor.sep_token_id()
    n_pos = model.encoder.embeddings.position_embeddings.num_embeddings
    mlf = lambda s: min(int(max_len_a * s + max_len_b), n_pos)            # :113-114
    if max_len is None:
        max_len = mlf(src_inputs.size(1))
    if src_inputs is None:
        max_lens = torch.LongTensor([max_len] * B).to(device)
    else:
        max_lens = torch.LongTensor([mlf(x) for x in src_sizes]).to(device)
    prefixes = first_tokens.unsqueeze(1).to(device)                       # (B,1)
    scores = torch.zeros(prefixes.size(), device=device)
    sizes = torch.zeros(B, device=device) if beam > 1 else None
    decoder = model.decoder[batch_lang] if model.lang_dec else model.decoder
    out_layer = model.output_layer if ((not model.lang_dec) and model.tie_embed) else model.output_layer[batch_lang]
    for i in range(1, max_len):
        cur = prefixes.view(-1, prefixes.size(-1))
        if int(torch.sum(torch.any(cur == eos, 1))) == beam * B:         # :134
            break
        rep = (lambda x: x) if i == 1 else (lambda x: torch.repeat_interleave(x, beam, 0))
        mem = rep(memory)
        types = rep(tgt_langs.unsqueeze(-1).expand(-1, cur.size(1)).to(device))
        mmask = rep(src_mask) if src_inputs is not None else None
        states = decoder(encoder_states=mem, input_ids=cur, encoder_attention_mask=mmask,
                         tgt_attention_mask=torch.ones(cur.size(), device=device), token_type_ids=types)
        if memory2 is not None:                                           # :167-188
            st2 = decoder2(encoder_states=rep(memory2.to(device)), input_ids=cur, encoder_attention_mask=None,
                           tgt_attention_mask=torch.ones(cur.size(), device=device), token_type_ids=types)
            sig = torch.sigmoid(gate + 1e-7)
            states = sig * states + (1 - sig) * st2
        last = states[:, -1, :]                                           # :190
        if model.use_proposals:                                           # :191-192
            last = model.attend_proposal(last, proposals, pad_idx)
        logp = F.log_softmax(out_layer(last), dim=-1)                     # :194
        prefixes, scores, sizes, raw_idx = search_step(i, logp, prefixes, scores, sizes, max_lens, eos, pad_idx,
                                                       beam, len_penalty_ratio, topk_fn)
        if trace is not None:
            trace.append(dict(i=i, last_state=last.clone(), log_probs=logp.clone(), prefixes=prefixes.clone(),
                              scores=scores.clone(), sizes=None if sizes is None else sizes.clone(),
                              raw_idx=raw_idx.clone()))
    outputs = prefixes[:, 0, :]                                           # :229
    if unpad_output:
        return outputs_until_eos(eos, outputs, size_limit=max_lens)
    outputs = outputs.cpu()
    return [outputs[r] for r in range(outputs.size(0))]
