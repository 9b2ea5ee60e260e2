"""Host side of the decode path: weight packing, workspace, and the step loop over the C ABI.

Data layout in HBM (all allocated once per engine and reused across batches; kernels never allocate):
  weights    per decoder layer: Wqkv [3d][d] (query|key|value stacked), Wo, Wcq, Wco [d][d], W1 [F][d], W2 [d][F]
             in the compute dtype; biases / LayerNorm parameters / embedding tables in fp32;
             cross K,V projections of ALL layers stacked into one [L*2d][d] matrix (one GEMM per batch).
  cross_kv   [B][S][L][2][d]      cross-attention K,V, computed once per batch (reference: every step)
  kv_self    [L][M][T][2][d]      self-attention KV cache, M = B*beam rows; NEVER reordered: the beam kernel
                                  maintains anc[M][T] (which beam slot physically holds position p)
  tokens/anc/scores/sizes/fin     ping-pong search state written by the fused beam kernel
  cand_*, chunk_*                 per (row, 128-entry vocabulary chunk) max / sum-exp / top-K written by the fused vocabulary
                                  GEMM epilogue (16-bit modes): no (M,V) tensor exists in HBM
  logits     [M][V4] fp32         fp32 parity mode and parity traces only (written once by the CUDA-core GEMM, read once)
"""
import ctypes as C

import torch

from . import _lib

_ACT = {_lib.F32: torch.float32, _lib.BF16: torch.bfloat16, _lib.F16: torch.float16}      # activations / KV cache
_WT = {_lib.F32: torch.float32, _lib.BF16: torch.bfloat16, _lib.F16: torch.float16}      # linear weights


def _lin(mod):
    return mod.weight.detach(), mod.bias.detach()


_UID = [0]


def _src_key(tensors):
    """What a packed copy depends on: in-place updates bump `_version`; `model.half()` / `.bfloat16()` / `.to(device)` swap
    `param.data` WITHOUT bumping it, so storage address, dtype and device are part of the key too."""
    return tuple((t._version, t.data_ptr(), t.dtype, str(t.device)) for t in tensors)


class PackedDecoder(object):
    """Device-resident weights of ONE decoder (+ its output layer) in kernel layout.

    Source parameter names are the reference's state-dict keys:
    decoder.embeddings.*, decoder.decoder.layer.N.{attention,crossattention}.{self.{query,key,value},output.dense,
    output.LayerNorm}, .intermediate.dense, .output.{dense,LayerNorm}; output_layer[.K].layer.{weight,bias}.
    """

    def __init__(self, decoder, output_layer, dtype_code, device):
        act = _WT[dtype_code]       # weight dtype of every linear
        self.dtype_code = dtype_code
        self._src = []
        _UID[0] += 1
        self.uid = _UID[0]           # cache key of everything that holds raw pointers into these weights

        def f32(t):
            self._src.append(t)
            return t.detach().to(device=device, dtype=torch.float32).contiguous()

        def mat(t):
            self._src.append(t)
            return t.detach().to(device=device, dtype=act).contiguous()

        emb = decoder.embeddings
        self.emb_word, self.emb_pos, self.emb_type = (f32(emb.word_embeddings.weight), f32(emb.position_embeddings.weight),
                                                      f32(emb.token_type_embeddings.weight))
        self.emb_ln_g, self.emb_ln_b = f32(emb.LayerNorm.weight), f32(emb.LayerNorm.bias)
        self.ln_eps = float(emb.LayerNorm.eps)
        self.n_pos = emb.position_embeddings.weight.shape[0]
        layers = decoder.decoder.layer
        self.L = len(layers)
        self.d = emb.word_embeddings.weight.shape[1]
        self.H = layers[0].attention.self.num_attention_heads
        self.F = layers[0].intermediate.dense.weight.shape[0]
        self.keep = []
        self.layers = (_lib.Layer * self.L)()
        ckv_w, ckv_b = [], []
        for i, ly in enumerate(layers):
            sa, ca = ly.attention, ly.crossattention
            for m in (sa.self.query, sa.self.key, sa.self.value, ca.self.key, ca.self.value):
                self._src.extend([m.weight, m.bias])
            wqkv = torch.cat([sa.self.query.weight, sa.self.key.weight, sa.self.value.weight], 0).detach().to(device, act).contiguous()
            bqkv = torch.cat([sa.self.query.bias, sa.self.key.bias, sa.self.value.bias], 0).detach().to(device, torch.float32).contiguous()
            ckv_w += [ca.self.key.weight.detach(), ca.self.value.weight.detach()]
            ckv_b += [ca.self.key.bias.detach(), ca.self.value.bias.detach()]
            t = dict(w_qkv=wqkv, b_qkv=bqkv,
                     w_o=mat(sa.output.dense.weight), b_o=f32(sa.output.dense.bias),
                     ln1_g=f32(sa.output.LayerNorm.weight), ln1_b=f32(sa.output.LayerNorm.bias),
                     w_cq=mat(ca.self.query.weight), b_cq=f32(ca.self.query.bias),
                     w_co=mat(ca.output.dense.weight), b_co=f32(ca.output.dense.bias),
                     ln2_g=f32(ca.output.LayerNorm.weight), ln2_b=f32(ca.output.LayerNorm.bias),
                     w_1=mat(ly.intermediate.dense.weight), b_1=f32(ly.intermediate.dense.bias),
                     w_2=mat(ly.output.dense.weight), b_2=f32(ly.output.dense.bias),
                     ln3_g=f32(ly.output.LayerNorm.weight), ln3_b=f32(ly.output.LayerNorm.bias))
            self.keep.append(t)
            for k, v in t.items():
                setattr(self.layers[i], k, v.data_ptr())
        self.w_ckv = torch.cat(ckv_w, 0).to(device, act).contiguous()          # [L*2d][d]
        self.b_ckv = torch.cat(ckv_b, 0).to(device, torch.float32).contiguous()
        if output_layer is not None:
            self.w_vocab, self.b_vocab = mat(output_layer.layer.weight), f32(output_layer.layer.bias)
            self.V = self.w_vocab.shape[0]
        else:                        # second-stream stack (obj_decoder): shares the first stream's vocabulary projection
            self.w_vocab = self.b_vocab = None
            self.V = emb.word_embeddings.weight.shape[0]
        self.versions = _src_key(self._src)

    def stale(self):
        return _src_key(self._src) != self.versions


class PackedEncoder(object):
    """Device-resident weights of the bidirectional text encoder in kernel layout (reference src/bert_seq2seq.py:94-144;
    state-dict names encoder.embeddings.*, encoder.encoder.layer.N.attention.{self.{query,key,value},output.dense,
    output.LayerNorm}, .intermediate.dense, .output.{dense,LayerNorm}).  key|value are stacked into one [2d][d] matrix so
    that their GEMM writes the [position][K|V] rows the attention kernels read."""

    def __init__(self, encoder, dtype_code, device):
        act = _WT[dtype_code]
        self.dtype_code = dtype_code
        self._src = []

        def f32(t):
            self._src.append(t)
            return t.detach().to(device=device, dtype=torch.float32).contiguous()

        def mat(t):
            self._src.append(t)
            return t.detach().to(device=device, dtype=act).contiguous()

        emb = encoder.embeddings
        self.emb_word, self.emb_pos, self.emb_type = (f32(emb.word_embeddings.weight), f32(emb.position_embeddings.weight),
                                                      f32(emb.token_type_embeddings.weight))
        self.emb_ln_g, self.emb_ln_b = f32(emb.LayerNorm.weight), f32(emb.LayerNorm.bias)
        self.ln_eps = float(emb.LayerNorm.eps)
        self.d = emb.word_embeddings.weight.shape[1]
        layers = encoder.encoder.layer
        self.H = layers[0].attention.self.num_attention_heads
        self.F = layers[0].intermediate.dense.weight.shape[0]
        self.layers = []
        for ly in layers:
            sa = ly.attention
            for m in (sa.self.key, sa.self.value):
                self._src.extend([m.weight, m.bias])
            self.layers.append(dict(
                w_q=mat(sa.self.query.weight), b_q=f32(sa.self.query.bias),
                w_kv=torch.cat([sa.self.key.weight, sa.self.value.weight], 0).detach().to(device, act).contiguous(),
                b_kv=torch.cat([sa.self.key.bias, sa.self.value.bias], 0).detach().to(device, torch.float32).contiguous(),
                w_o=mat(sa.output.dense.weight), b_o=f32(sa.output.dense.bias),
                ln1_g=f32(sa.output.LayerNorm.weight), ln1_b=f32(sa.output.LayerNorm.bias),
                w_1=mat(ly.intermediate.dense.weight), b_1=f32(ly.intermediate.dense.bias),
                w_2=mat(ly.output.dense.weight), b_2=f32(ly.output.dense.bias),
                ln2_g=f32(ly.output.LayerNorm.weight), ln2_b=f32(ly.output.LayerNorm.bias)))
        self.versions = _src_key(self._src)

    def stale(self):
        return _src_key(self._src) != self.versions


class Workspace(object):
    """Flat device buffers that only ever grow; indexing always uses the CURRENT batch shape."""

    def __init__(self, device):
        self.device = device
        self.buf = {}
        self.generation = 0          # bumped whenever a buffer moves (cached decoders / graphs hold raw pointers)

    def get(self, name, numel, dtype):
        t = self.buf.get(name)
        if t is None or t.numel() < numel or t.dtype != dtype:
            # grow geometrically: every growth invalidates the cached decoders (raw pointers), so it has to be rare
            grown = int(numel) if t is None or t.dtype != dtype else max(int(numel), int(t.numel() * 1.5))
            t = torch.empty(max(grown, 1), dtype=dtype, device=self.device)
            self.buf[name] = t
            self.generation += 1
        return t


class DecodeEngine(object):
    GRAPH_STEPS = 8          # search steps per CUDA-graph replay
    MAX_CACHED = 256         # decoders (TMA descriptors + instantiated graph) kept per engine, least recently used first out
    S_QUANTUM, T_QUANTUM = 8, 16     # memory length / cache capacity are rounded up to these, so that a corpus of ragged
                                     # batches maps onto a small lattice of decoder shapes (the real lengths travel as data:
                                     # mem_len per sentence, loop bound per group)

    def __init__(self, device, dtype_code, gemm_impl=_lib.GEMM_AUTO, poll_every=16, use_graph=True):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("imagetranslate_b200: the decode path runs on CUDA devices only (no CPU fallback)")
        self.dtype_code = dtype_code
        self.gemm_impl = gemm_impl
        self.poll_every = poll_every
        self.ws = Workspace(self.device)
        self.lib = _lib.lib()
        self.use_graph = use_graph
        self.last_steps = 0
        self.launches_per_step = None
        self._decoders = {}          # shape key -> itb_decoder* (C.c_void_p)
        self.decoders_created = 0
        self._dec_generation = -1
        self._ctl_host = [torch.empty(8, dtype=torch.int32).pin_memory() for _ in range(4)]

    def _drop_decoders(self):
        for dec in self._decoders.values():
            self.lib.itb_decoder_destroy(dec)
        self._decoders.clear()

    def __del__(self):
        try:
            self._drop_decoders()
        except Exception:
            pass

    def _decoder(self, key, cfg):
        """itb_decoder for this (weights, shape): cached, because creating one encodes ~40 TMA descriptors and
        capturing its graph costs about as much as a few search steps."""
        if self._dec_generation != self.ws.generation:       # some workspace buffer moved: every cached pointer is stale
            self._drop_decoders()
            self._dec_generation = self.ws.generation
        dec = self._decoders.pop(key, None)
        if dec is None:
            if len(self._decoders) >= self.MAX_CACHED:
                self.lib.itb_decoder_destroy(self._decoders.pop(next(iter(self._decoders))))
            dec = C.c_void_p()
            _lib.check(self.lib.itb_decoder_create(C.byref(cfg), C.byref(dec)), "decoder_create")
            self.decoders_created += 1
        self._decoders[key] = dec           # (re-)inserted last: dict order = recency
        return dec

    def _config(self, pw, B, beam, S, T_max, pad, eos, alpha, G=1, want_logits=False, extra=None, S2=0, n_prop=0):
        ws, act, M = self.ws, _ACT[self.dtype_code], B * beam
        V4 = (pw.V + 3) // 4 * 4
        # 16-bit tensor-core decoders: the vocabulary GEMM's epilogue emits the per-chunk max / sum-exp / top-K itself, no
        # (M,V) logits tensor exists (it is only materialised for the parity traces of the tests)
        fused = self.dtype_code != _lib.F32 and self.gemm_impl != _lib.GEMM_SIMT
        if fused:
            n_chunk = self.lib.itb_vocab_chunks(pw.V)
        else:
            n_chunk = max(1, min(32, -(-2368 // M), pw.V // 1024 if pw.V >= 1024 else 1))
        KT = self.lib.itb_topk_capacity(beam)
        c = _lib.Config()
        c.dtype, c.gemm_impl = self.dtype_code, self.gemm_impl
        c.B, c.beam, c.d, c.H, c.F, c.L, c.V, c.S, c.T_max = B, beam, pw.d, pw.H, pw.F, pw.L, pw.V, S, T_max
        c.pad_id, c.eos_id, c.n_chunk, c.G, c.fused_vocab = pad, eos, n_chunk, G, int(fused)
        c.len_alpha, c.ln_eps = alpha, pw.ln_eps
        c.emb_word, c.emb_pos, c.emb_type = pw.emb_word.data_ptr(), pw.emb_pos.data_ptr(), pw.emb_type.data_ptr()
        c.emb_ln_g, c.emb_ln_b = pw.emb_ln_g.data_ptr(), pw.emb_ln_b.data_ptr()
        c.layers_host = C.cast(pw.layers, C.c_void_p)
        c.w_vocab, c.b_vocab = pw.w_vocab.data_ptr(), pw.b_vocab.data_ptr()
        i32, u8, f32 = torch.int32, torch.uint8, torch.float32
        t = dict(
            tgt_lang=ws.get("tgt_lang", B, i32), max_lens=ws.get("max_lens", B, i32),
            sent_grp=ws.get("sent_grp", B, i32), mem_len=ws.get("mem_len", B, i32), grp=ws.get("grp", 8 * G, i32),
            sent_done=ws.get("sent_done", B, i32),
            src_mask=ws.get("src_mask", B * S, u8), cross_kv=ws.get("cross_kv", B * S * pw.L * 2 * pw.d, act),
            ctl=ws.get("ctl", 8, i32), tokens=ws.get("tokens", 2 * M * T_max, i32), anc=ws.get("anc", 2 * M * T_max, u8),
            scores=ws.get("scores", 2 * M, f32), sizes=ws.get("sizes", 2 * M, f32), fin=ws.get("fin", 2 * M, u8),
            kv_self=ws.get("kv_self", pw.L * M * T_max * 2 * pw.d, act),
            x_f32=ws.get("x_f32", M * pw.d, f32), x_act=ws.get("x_act", M * pw.d, act), qkv=ws.get("qkv", M * 3 * pw.d, f32),
            ctx=ws.get("ctx", M * pw.d, act), y=ws.get("y", _lib.Y_PARTIALS * M * pw.d, f32), ffn_h=ws.get("ffn_h", M * pw.F, act),
            logits=ws.get("logits", M * V4 if (want_logits or not fused) else 1, f32), cand_val=ws.get("cand_val", M * n_chunk * KT, f32),
            cand_idx=ws.get("cand_idx", M * n_chunk * KT, i32), chunk_max=ws.get("chunk_max", M * n_chunk, f32),
            chunk_sum=ws.get("chunk_sum", M * n_chunk, f32), sent_counter=ws.get("sent_counter", B, i32))
        for k, v in t.items():
            setattr(c, k, v.data_ptr())
        if fused and not want_logits:
            c.logits = None
        t["out_tok"], t["out_len"] = ws.get("out_tok", B * T_max, i32), ws.get("out_len", 2 * B + 8, i32)
        t["first"], t["grp_max_len"] = ws.get("first", B, i32), ws.get("grp_max_len", G, i32)
        t["memory"] = ws.get("memory", B * S * pw.d, act)
        extra = extra or {}
        if extra.get("pw2") is not None:          # second decoder stream (reference src/seq_gen.py:167-188)
            pw2 = extra["pw2"]
            if (pw2.L, pw2.d, pw2.H, pw2.F) != (pw.L, pw.d, pw.H, pw.F):
                raise RuntimeError("the second decoder stream must have the first one's shape (layers, hidden, heads, FFN)")
            c.two_stream, c.S2 = 1, S2
            c.layers2_host = C.cast(pw2.layers, C.c_void_p)
            c.emb2_word, c.emb2_pos, c.emb2_type = pw2.emb_word.data_ptr(), pw2.emb_pos.data_ptr(), pw2.emb_type.data_ptr()
            c.emb2_ln_g, c.emb2_ln_b = pw2.emb_ln_g.data_ptr(), pw2.emb_ln_b.data_ptr()
            t["memory2"] = ws.get("memory2", B * S2 * pw.d, act)
            t["cross_kv2"] = ws.get("cross_kv2", B * S2 * pw.L * 2 * pw.d, act)
            t["kv_self2"] = ws.get("kv_self2", pw.L * M * T_max * 2 * pw.d, act)
            t["x_first"] = ws.get("x_first", M * pw.d, f32)
            t["mem_len2"] = ws.get("mem_len2", B, i32)
            c.cross_kv2, c.kv_self2, c.x_first = t["cross_kv2"].data_ptr(), t["kv_self2"].data_ptr(), t["x_first"].data_ptr()
            c.mem_len2, c.src_mask2 = t["mem_len2"].data_ptr(), None
            c.stream_gate, c.gate_eps = extra["gate"].data_ptr(), float(extra.get("gate_eps", 1e-7))
        if n_prop > 0:                            # lexical proposals (reference src/seq2seq.py:110-144)
            lex = extra["lex"]
            t["proposals"] = ws.get("proposals", B * n_prop, i32)
            c.n_prop, c.proposals = n_prop, t["proposals"].data_ptr()
            c.prop_emb, c.lex_gate = lex["emb"].data_ptr(), lex["gate"].data_ptr()
            c.lex_ln_g, c.lex_ln_b, c.lex_ln_eps = lex["ln_g"].data_ptr(), lex["ln_b"].data_ptr(), float(lex["eps"])
        return c, t

    @torch.no_grad()
    def encode(self, pe, src_inputs, src_mask, src_langs):
        """Encoder forward of one padded source batch on the C-ABI kernels (replaces reference src/seq2seq.py:101-108 ->
        src/bert_seq2seq.py:103-144, HF BertLayer without cross-attention: Appendix A.3-A.6): embeddings+LN, then per layer
        Q / K|V projections, bidirectional attention with the additive -10000 padding mask (every position of a sentence is
        a query over the sentence's K,V rows: the single-query kernels with beam := S), output projection + residual + LN,
        FFN1 + erf-GELU, FFN2 + residual + LN.  src_inputs (B,S) ints, src_mask (B,S) bool, src_langs (B,) ints.
        Returns the memory (B,S,d) in the activation dtype (a fresh tensor)."""
        lib, dev, ws, code = self.lib, self.device, self.ws, self.dtype_code
        act, f32, i32 = _ACT[code], torch.float32, torch.int32
        B, S = int(src_inputs.shape[0]), int(src_inputs.shape[1])
        d, H, F, rows = pe.d, pe.H, pe.F, B * S
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        tok = src_inputs.to(device=dev, dtype=i32, non_blocking=True).contiguous()
        lang = src_langs.to(device=dev, dtype=i32, non_blocking=True).contiguous()
        mask = src_mask.to(device=dev, non_blocking=True).to(torch.uint8).contiguous()
        x_f32, x_act = ws.get("enc_x_f32", rows * d, f32), ws.get("enc_x_act", rows * d, act)
        q, kv = ws.get("enc_q", rows * d, f32), ws.get("enc_kv", rows * 2 * d, act)
        ctx, y, h = ws.get("enc_ctx", rows * d, act), ws.get("enc_y", rows * d, f32), ws.get("enc_h", rows * F, act)
        P, impl = _lib.ptr, self.gemm_impl
        _lib.check(lib.itb_embed_seq_ln(code, P(tok), P(lang), P(pe.emb_word), P(pe.emb_pos), P(pe.emb_type), P(pe.emb_ln_g),
                                        P(pe.emb_ln_b), pe.ln_eps, P(x_f32), P(x_act), B, S, d, stream), "embed_seq_ln")
        for w in pe.layers:
            _lib.check(lib.itb_linear(code, impl, _lib.EPI_BIAS_F32, P(x_act), P(w["w_q"]), P(w["b_q"]), P(q), rows, d, d, d, stream), "enc q")
            _lib.check(lib.itb_linear(code, impl, _lib.EPI_BIAS_ACT, P(x_act), P(w["w_kv"]), P(w["b_kv"]), P(kv), rows, 2 * d, d, 2 * d,
                                      stream), "enc kv")
            _lib.check(lib.itb_cross_attention(code, None, None, None, P(q), d, P(kv), 0, P(mask), P(ctx), rows, S, d, H, S, 1, stream),
                       "enc attention")
            _lib.check(lib.itb_linear(code, impl, _lib.EPI_BIAS_F32, P(ctx), P(w["w_o"]), P(w["b_o"]), P(y), rows, d, d, d, stream), "enc o")
            _lib.check(lib.itb_add_ln(code, None, P(y), 1, P(w["ln1_g"]), P(w["ln1_b"]), pe.ln_eps, P(x_f32), P(x_act), rows, d, stream), "enc ln1")
            _lib.check(lib.itb_linear(code, impl, _lib.EPI_BIAS_GELU_ACT, P(x_act), P(w["w_1"]), P(w["b_1"]), P(h), rows, F, d, F, stream), "enc ffn1")
            _lib.check(lib.itb_linear(code, impl, _lib.EPI_BIAS_F32, P(h), P(w["w_2"]), P(w["b_2"]), P(y), rows, d, F, d, stream), "enc ffn2")
            _lib.check(lib.itb_add_ln(code, None, P(y), 1, P(w["ln2_g"]), P(w["ln2_b"]), pe.ln_eps, P(x_f32), P(x_act), rows, d, stream), "enc ln2")
        self.enc_launches = 1 + 8 * len(pe.layers)
        return x_act[:rows * d].view(B, S, d).clone()

    @torch.no_grad()
    def decode(self, pw, memory, src_mask, first_tokens, tgt_type_ids, max_lens, max_len, beam, pad, eos, alpha, trace=None):
        """One reference batch.  memory (B,S,d) on device; src_mask (B,S) bool or None; first_tokens/tgt_type_ids/
        max_lens (B,) ints.  Returns (tokens (B,T_max) int32 CPU, lengths (B,), full_lengths (B,))."""
        return self.decode_groups(pw, [dict(memory=memory, src_mask=src_mask, first_tokens=first_tokens,
                                            tgt_type_ids=tgt_type_ids, max_lens=max_lens, max_len=max_len)],
                                  beam, pad, eos, alpha, trace)[0]

    @torch.no_grad()
    def decode_groups(self, pw, groups, beam, pad, eos, alpha, trace=None, extra=None):
        """Decodes several reference batches ("groups") in lockstep as ONE super-batch: their rows share every kernel
        launch and every weight byte streamed per step, while each group keeps its own loop bound, termination rule and
        memory length, so each group's result is exactly what decoding it alone gives.  groups: list of dicts with the
        arguments of `decode`.  Returns a list of (tokens (B_g,T_max) int32 CPU, lengths, full_lengths)."""
        lib, dev = self.lib, self.device
        G = len(groups)
        # Row layout: groups in order of decreasing loop bound, so that the rows of groups whose search has ended form a
        # SUFFIX of the row range -- the beam kernel publishes the live prefix (ctl.rows_active) and the GEMM / LayerNorm
        # kernels skip the rest.  A group's arithmetic does not depend on where its rows sit; results go back in call order.
        order = sorted(range(G), key=lambda i: (-int(groups[i]["max_len"]), i))
        groups = [groups[i] for i in order]
        Bs = [int(g["memory"].shape[0]) for g in groups]
        Ss = [int(g["memory"].shape[1]) for g in groups]
        q = lambda v, m: -(-int(v) // m) * m
        B, S, d = sum(Bs), q(max(Ss), self.S_QUANTUM), int(groups[0]["memory"].shape[2])
        if d != pw.d:
            raise RuntimeError("memory width %d does not match the decoder width %d" % (d, pw.d))
        max_len = max(int(g["max_len"]) for g in groups)
        T_max = q(max(max_len, 2), self.T_QUANTUM)
        has_mask = any(g["src_mask"] is not None for g in groups)
        extra = extra or {}
        two = extra.get("pw2") is not None
        S2s = [int(g["memory2"].shape[1]) for g in groups] if two else []
        S2 = q(max(S2s), self.S_QUANTUM) if two else 0
        props = [g.get("proposals") for g in groups]
        n_prop = max(int(p.shape[1]) for p in props) if extra.get("lex") is not None else 0
        cfg, t = self._config(pw, B, beam, S, T_max, pad, eos, alpha, G, want_logits=trace is not None, extra=extra, S2=S2,
                              n_prop=n_prop)
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        i32 = torch.int32
        cat = lambda k: torch.cat([g[k].to(device=dev, dtype=i32, non_blocking=True).reshape(-1) for g in groups])
        t["tgt_lang"][:B].copy_(cat("tgt_type_ids"), non_blocking=True)
        t["max_lens"][:B].copy_(cat("max_lens"), non_blocking=True)
        t["first"][:B].copy_(cat("first_tokens"), non_blocking=True)
        meta = torch.tensor([gi for gi, n in enumerate(Bs) for _ in range(n)] + [Ss[gi] for gi, n in enumerate(Bs) for _ in range(n)]
                            + [int(g["max_len"]) for g in groups]
                            + ([S2s[gi] for gi, n in enumerate(Bs) for _ in range(n)] if two else []),
                            dtype=i32).pin_memory().to(dev, non_blocking=True)
        t["sent_grp"][:B].copy_(meta[:B], non_blocking=True)
        t["mem_len"][:B].copy_(meta[B:2 * B], non_blocking=True)
        t["grp_max_len"][:G].copy_(meta[2 * B:2 * B + G], non_blocking=True)
        if two:                                   # second memory of every group, zero-padded to the longest
            t["mem_len2"][:B].copy_(meta[2 * B + G:], non_blocking=True)
            mem2_all = t["memory2"][:B * S2 * d].view(B, S2, d)
            if any(Sg != S2 for Sg in S2s):
                mem2_all.zero_()
            o = 0
            for g, Bg, Sg in zip(groups, Bs, S2s):
                mem2_all[o:o + Bg, :Sg].copy_(g["memory2"], non_blocking=True)
                o += Bg
            pw2 = extra["pw2"]
            _lib.check(lib.itb_cross_kv(self.dtype_code, self.gemm_impl, _lib.ptr(mem2_all), _lib.ptr(pw2.w_ckv), _lib.ptr(pw2.b_ckv),
                                        _lib.ptr(t["cross_kv2"]), B * S2, pw.d, pw.L, stream), "cross_kv2")
        if n_prop > 0:                            # proposals of every group, PAD-padded to the widest
            pr_all = t["proposals"][:B * n_prop].view(B, n_prop)
            pr_all.fill_(pad)
            o = 0
            for p_, Bg in zip(props, Bs):
                pr_all[o:o + Bg, :p_.shape[1]].copy_(p_.to(device=dev, dtype=i32), non_blocking=True)
                o += Bg
        mem_all = t["memory"][:B * S * d].view(B, S, d)
        mask_all = t["src_mask"][:B * S].view(B, S)
        if any(Sg != S for Sg in Ss):       # padded positions are never attended (mem_len), but their K,V must be finite
            mem_all.zero_()
        o = 0
        for g, Bg, Sg in zip(groups, Bs, Ss):
            mem_all[o:o + Bg, :Sg].copy_(g["memory"], non_blocking=True)
            if has_mask:
                if g["src_mask"] is not None:
                    mask_all[o:o + Bg, :Sg].copy_(g["src_mask"].to(device=dev), non_blocking=True)
                else:
                    mask_all[o:o + Bg, :Sg].fill_(1)
            o += Bg
        if not has_mask:
            cfg.src_mask = None
        _lib.check(lib.itb_cross_kv(self.dtype_code, self.gemm_impl, _lib.ptr(mem_all), _lib.ptr(pw.w_ckv), _lib.ptr(pw.b_ckv),
                                    _lib.ptr(t["cross_kv"]), B * S, pw.d, pw.L, stream), "cross_kv")
        _lib.check(lib.itb_init_state(C.byref(cfg), _lib.ptr(t["first"]), _lib.ptr(t["grp_max_len"]), stream), "init_state")
        src_mask = True if has_mask else None
        key = (pw.uid, B, beam, S, T_max, has_mask, pad, eos, float(alpha), G, trace is not None,
               extra["pw2"].uid if two else 0, S2, n_prop)
        dec = self._decoder(key, cfg)
        n_steps = max(int(max_len) - 1, 0)
        if trace is None and self.use_graph and n_steps > 0:
            # CUDA-graph replay, GRAPH_STEPS search steps per launch; the device-side ctl block carries the step index
            # and the termination flag, so the host only peeks at ctl.done (asynchronously, two replays behind).
            _lib.check(lib.itb_decoder_capture(dec, self.GRAPH_STEPS), "decoder_capture")
            n_launch = -(-n_steps // self.GRAPH_STEPS)
            pending = []
            for i in range(n_launch):
                _lib.check(lib.itb_decoder_launch(dec, stream), "decoder_launch")
                if n_launch - i > 3:
                    ev = torch.cuda.Event()
                    hb = self._ctl_host[i % 4]
                    hb.copy_(t["ctl"][:8], non_blocking=True)
                    ev.record()
                    pending.append((ev, hb))
                    if len(pending) > 2:
                        ev0, hb0 = pending.pop(0)
                        ev0.synchronize()
                        if int(hb0[1]) != 0:
                            break
        else:
            i = 0
            while i < n_steps:
                if trace is not None:      # test hook: state BEFORE the step (forces a sync per step)
                    ctl = t["ctl"][:8].cpu()
                    if int(ctl[1]) != 0:
                        break
                    M, V4 = B * beam, (pw.V + 3) // 4 * 4
                    cur, step = int(ctl[3]), int(ctl[0])
                    pre = t["tokens"][:2 * M * T_max].view(2, M, T_max)[cur, :, :step].clone().cpu()
                _lib.check(lib.itb_decoder_step(dec, stream), "decoder_step")
                if trace is not None:
                    M, V4 = B * beam, (pw.V + 3) // 4 * 4
                    ctl2 = t["ctl"][:8].cpu()
                    trace.append(dict(step=step, prefixes=pre,
                                      logits=t["logits"][:M * V4].view(M, V4)[:, :pw.V].clone().cpu(),
                                      scores=t["scores"][:2 * M].view(2, M)[int(ctl2[3])].clone().cpu(),
                                      tokens=t["tokens"][:2 * M * T_max].view(2, M, T_max)[int(ctl2[3]), :, :step + 1].clone().cpu()))
                i += 1
                if i % self.poll_every == 0 and i < n_steps and int(t["ctl"][1].item()) != 0:   # ctl.done
                    break
        out_tok, out_len = t["out_tok"], t["out_len"]
        _lib.check(lib.itb_finalize(_lib.ptr(t["ctl"]), _lib.ptr(t["grp"]), _lib.ptr(t["sent_grp"]), _lib.ptr(t["tokens"]),
                                    _lib.ptr(t["max_lens"]), eos,
                                    _lib.ptr(out_tok), _lib.ptr(out_len), C.c_void_p(out_len.data_ptr() + 4 * B),
                                    B, beam, T_max, stream), "finalize")
        tok = out_tok[:B * T_max].view(B, T_max).cpu()
        out_len[2 * B:2 * B + 8].copy_(t["ctl"][:8])
        lens = out_len[:2 * B + 8].cpu()
        self.last_steps = int(lens[2 * B + 2])
        self.launches_per_step = 2 + 11 * pw.L + ((3 + 11 * pw.L) if two else 0) + (1 if n_prop else 0)
        res, o = [None] * G, 0
        for gi, Bg in zip(order, Bs):
            res[gi] = (tok[o:o + Bg], lens[o:o + Bg], lens[B + o:B + o + Bg])
            o += Bg
        return res
