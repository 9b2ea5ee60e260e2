"""Drop-in `BeamDecoder` (reference src/seq_gen.py:27-242) over the B200 decode engine.

Same constructor and `forward` signature, same list-unwrapping of DataParallel-style kwargs
(:57-72), same max_len / max_lens rules (:113-121), same return value (list of 1-D CPU LongTensors
that start with the first/language token and exclude EOS; `unpad_output=False` returns full rows).
The search itself (decoder step, vocabulary projection, log-softmax, scoring, top-k, back-pointers,
termination) runs in CUDA kernels -- see include/itb.h.  The encoder memory still comes from the
model's own `encode` (:96,99,106), exactly as in the reference.
"""
import os
import threading
from concurrent.futures import ThreadPoolExecutor

import torch
import torch.nn as nn

from . import _lib
from .engine import DecodeEngine, PackedDecoder, PackedEncoder


def get_outputs_until_eos(eos, outputs, size_limit=None, remove_first_token: bool = False):
    """reference src/seq_gen.py:6-24: rows cut before their first EOS (else at size_limit[r])."""
    if outputs.dim() == 1:
        outputs = outputs.unsqueeze(0)
    outputs = outputs.cpu()
    start = 1 if remove_first_token else 0
    res = []
    for r in range(outputs.size(0)):
        hits = torch.nonzero(outputs[r] == eos)
        if hits.numel() > 0:
            res.append(outputs[r, start:int(hits[0, 0])])
        else:
            last = outputs.size(1) if size_limit is None else int(size_limit[r])
            res.append(outputs[r, start:last])
    return res


class BeamDecoder(nn.Module):
    def __init__(self, seq2seq_model, beam_width: int = 5, max_len_a: float = 1.1, max_len_b: int = 5,
                 len_penalty_ratio: float = 0.8, dtype: str = "auto", gemm_impl: int = _lib.GEMM_AUTO):
        super(BeamDecoder, self).__init__()
        self.seq2seq_model = seq2seq_model
        self.beam_width = beam_width
        self.max_len_a = max_len_a
        self.max_len_b = max_len_b
        self.len_penalty_ratio = len_penalty_ratio
        self.compute_dtype = dtype          # "auto": fp32 weights -> fp32 kernels, half/bf16 weights -> bf16 tensor cores
        self.gemm_impl = gemm_impl
        self._engines = {}
        self._packed = {}
        self._lane = threading.local()      # forward_many: which engine (workspace + stream) this thread drives
        self._lock = threading.Lock()
        self.kernel_launches = 0            # kernels enqueued by this decoder so far (bench.py's gpu_launches)
        self.trace = None                   # parity hook: a list here receives the per-step state (one host sync per step)
        # forward_many packs consecutive batches into one super-batch while their memory lengths stay within this ratio
        # (rows of the shorter batches idle once their own search has ended)
        self.pack_ratio = float(os.environ.get("ITB_PACK_RATIO", "1.5"))
        # text memory: run the encoder on this library's kernels (False / ITB_TORCH_ENCODER=1: call model.encode instead)
        self.cuda_encoder = os.environ.get("ITB_TORCH_ENCODER", "0") != "1"

    def len_penalty(self, lengths: torch.Tensor):
        """reference src/seq_gen.py:37-44 (kept for callers; the kernels compute it on device)."""
        return torch.pow((lengths + 6.0) / 6.0, self.len_penalty_ratio).unsqueeze(-1)

    # ------------------------------------------------------------------ helpers
    def _dtype_code(self, weight):
        if self.compute_dtype in ("fp32", "float32", "f32"):
            return _lib.F32
        if self.compute_dtype in ("bf16", "bfloat16"):
            return _lib.BF16
        if self.compute_dtype in ("fp16", "float16", "half"):
            return _lib.F16            # fp16 weights, activations and KV cache (the reference's --fp16 / apex O2 arithmetic)
        return _lib.F32 if weight.dtype == torch.float32 else _lib.F16

    def _engine(self, device, code):
        key = (str(device), code, getattr(self._lane, "idx", 0))
        with self._lock:
            if key not in self._engines:
                self._engines[key] = DecodeEngine(device, code, self.gemm_impl)
            return self._engines[key]

    def _weights(self, model, lang, code, device):
        key = (lang if model.lang_dec or not model.tie_embed else -1, code, str(device))
        with self._lock:
            pw = self._packed.get(key)
        if pw is None or pw.stale():
            decoder = model.decoder[lang] if model.lang_dec else model.decoder
            out_layer = model.output_layer if ((not model.lang_dec) and model.tie_embed) else model.output_layer[lang]
            pw = PackedDecoder(decoder, out_layer, code, device)
            self._packed[key] = pw
        return pw

    def _stack_weights(self, key, decoder, code, device):
        """Packed weights of a decoder stack that has no output layer of its own (obj_decoder: src/image_model.py:279-296)."""
        key = (key, code, str(device))
        with self._lock:
            pw = self._packed.get(key)
        if pw is None or pw.stale():
            pw = PackedDecoder(decoder, None, code, device)
            self._packed[key] = pw
        return pw

    def _enc_weights(self, model, code, device):
        key = ("enc", code, str(device))
        with self._lock:
            pe = self._packed.get(key)
        if pe is None or pe.stale():
            pe = PackedEncoder(model.encoder, code, device)
            self._packed[key] = pe
        return pe

    @staticmethod
    def _is_bert_text_encoder(model):
        enc = getattr(model, "encoder", None)
        return (enc is not None and hasattr(enc, "embeddings") and hasattr(getattr(enc, "encoder", None), "layer")
                and all(hasattr(ly, "attention") and hasattr(ly, "intermediate") and not hasattr(ly, "crossattention")
                        for ly in enc.encoder.layer))

    def refresh_weights(self):
        """Drop packed weights (call after loading a new checkpoint into the wrapped model)."""
        self._packed.clear()

    # ------------------------------------------------------------------ forward
    def _prepare(self, src_inputs=None, src_sizes=None, first_tokens=None, src_mask=None, src_langs=None, tgt_langs=None,
                 pad_idx=None, max_len: int = None, unpad_output: bool = True, beam_width: int = None, images=None,
                 proposals=None, image_embed=None):
        """Everything of reference src/seq_gen.py:57-129 that happens before the search loop: argument unwrapping,
        the encoder call, max_len / max_lens.  Returns one decode "group" (= this reference batch)."""
        if isinstance(tgt_langs, list):                       # :57-72
            assert len(tgt_langs) == 1
            tgt_langs = tgt_langs[0]
            first_tokens = first_tokens[0]
        if isinstance(src_langs, list):
            src_langs, src_mask, src_sizes, src_inputs = src_langs[0], src_mask[0], src_sizes[0], src_inputs[0]
        if isinstance(images, list):
            images = images[0]
        if isinstance(image_embed, list):
            image_embed = image_embed[0]
        if isinstance(proposals, list):
            proposals = proposals[0]
        if beam_width is None:
            beam_width = self.beam_width
        model = self.seq2seq_model.module if hasattr(self.seq2seq_model, "module") else self.seq2seq_model
        wemb = model.encoder.embeddings.word_embeddings.weight
        device = wemb.device
        if device.type != "cuda":
            raise RuntimeError("imagetranslate_b200.BeamDecoder needs the model on a CUDA device (no CPU fallback)")
        use_prop = bool(getattr(model, "use_proposals", False))
        if use_prop and proposals is None:
            raise RuntimeError("the model uses lexical proposals (use_proposals) but no `proposals` were given")
        batch_lang = int(tgt_langs[0])
        memory2 = stream2 = None             # second decoder stream (:167-188)

        if images is not None and max_len is None:
            max_len = 512                                         # :90-91
        with torch.no_grad():
            if src_inputs is not None:                            # :94-96 / :105-106
                src_mask = src_mask.to(device)
                code0 = self._dtype_code(wemb)
                if self.cuda_encoder and self._is_bert_text_encoder(model):
                    # :96 Seq2Seq.encode -> BertEncoderModel.forward on this library's kernels (engine.encode)
                    eng = self._engine(device, code0)
                    memory = eng.encode(self._enc_weights(model, code0, device), src_inputs, src_mask, src_langs)
                    with self._lock:
                        self.kernel_launches += eng.enc_launches
                else:
                    langs2d = src_langs.unsqueeze(-1).expand(-1, src_inputs.size(-1))
                    memory = model.encode(src_inputs, src_mask, langs2d)[0]
                mem_mask = src_mask
                if images is not None:
                    # :105-106,180-188 text + image: the SAME decoder runs a second pass over the image memory, mixed by
                    # sigmoid(multimodal_attention_gate).  (As written, the reference's branch cannot run: its encode returns
                    # the image model's (grid, object) TUPLE as `image_embeddings`, 1-D src_langs, and never repeats the image
                    # memory per beam -- DESIGN.md §8.  Implemented here as the training forward of the same model defines it:
                    # src/image_model.py:205-216.)
                    img = model.image_model(images.to(device).to(wemb.dtype) if wemb.dtype != torch.float32 else images.to(device))
                    memory2 = img[0] if isinstance(img, (tuple, list)) else img
                    stream2 = ("same", model.multimodal_attention_gate)
            else:
                if image_embed is None:                           # :97-99
                    images = images.to(device)
                    if wemb.dtype == torch.float16:
                        images = images.half()
                    memory, obj_feat = model.encode(images=images)
                    if obj_feat is not None:                      # :167-179 object-feature stream through obj_decoder
                        memory2 = obj_feat
                        stream2 = ("obj", model.multistream_attention_gate)
                else:                                             # :100-103
                    memory = image_embed.to(device)
                mem_mask = None
        B = memory.size(0)
        eos = model.text_processor.sep_token_id()
        n_pos = model.encoder.embeddings.position_embeddings.num_embeddings
        mlf = lambda s: min(int(self.max_len_a * s + self.max_len_b), n_pos)   # :113-114
        if max_len is None:
            max_len = mlf(src_inputs.size(1))
        if src_inputs is None:
            max_lens = torch.full((B,), int(max_len), dtype=torch.int64)
        else:
            # :121 maps max_len_func over the ELEMENTS of the src_sizes LongTensor: `max_len_a * <0-dim int tensor>` is a
            # float32 tensor (python-scalar promotion), `+ max_len_b` stays float32, int() truncates.  float32, not double:
            # at a = 1.3 the two differ for s = 90, 170, 180, ... (float32 gives 121 for s = 90, double 122).
            max_lens = (src_sizes.to(torch.float32) * self.max_len_a + self.max_len_b).to(torch.int64).clamp(max=n_pos)
        max_len = min(int(max_len), n_pos)
        code = self._dtype_code(wemb)
        return dict(memory=memory, src_mask=mem_mask, first_tokens=first_tokens, tgt_type_ids=tgt_langs, max_lens=max_lens,
                    max_len=max_len, B=B, beam=int(beam_width), pad=int(pad_idx), eos=int(eos), unpad=bool(unpad_output),
                    lang=batch_lang, code=code, device=device, model=model, memory2=memory2, stream2=stream2,
                    proposals=proposals if use_prop else None)

    def _decode_groups(self, groups):
        """Runs one super-batch (groups sharing decoder weights / beam / pad / eos) and returns the reference's output
        lists (src/seq_gen.py:229-242), one per group."""
        g0 = groups[0]
        model, code, device = g0["model"], g0["code"], g0["device"]
        eng = self._engine(device, code)
        pw = self._weights(model, g0["lang"], code, device)
        extra = {}
        kinds = set(None if g["stream2"] is None else g["stream2"][0] for g in groups)
        if len(kinds) != 1 or len(set(g["proposals"] is None for g in groups)) != 1:
            raise RuntimeError("batches decoded in lockstep must all use the same decode variant")
        if g0["stream2"] is not None:
            kind, gate = g0["stream2"]
            if kind == "obj":
                obj_dec = model.obj_decoder[g0["lang"]] if model.lang_dec else model.obj_decoder
                extra["pw2"] = self._stack_weights(("obj", g0["lang"] if model.lang_dec else -1), obj_dec, code, device)
            else:
                extra["pw2"] = pw
            extra["gate"] = gate.detach().to(device=device, dtype=torch.float32).reshape(-1).contiguous()
            extra["gate_eps"] = 1e-7                             # :175,185
        if g0["proposals"] is not None:                          # :191-192 -> src/seq2seq.py:110-144
            f = lambda t: t.detach().to(device=device, dtype=torch.float32).contiguous()
            ln = model.lexical_layer_norm
            extra["lex"] = dict(emb=f(model.proposal_embedding.weight), gate=f(model.lexical_gate).reshape(-1), ln_g=f(ln.weight),
                                ln_b=f(ln.bias), eps=float(ln.eps))
        res = eng.decode_groups(pw, groups, g0["beam"], g0["pad"], g0["eos"], float(self.len_penalty_ratio), trace=self.trace,
                                extra=extra)
        with self._lock:
            self.kernel_launches += eng.last_steps * eng.launches_per_step + 4
        outs = []
        for g, (tok, lens, lens_full) in zip(groups, res):
            tok = tok.to(torch.int64)
            sel = (lens if g["unpad"] else lens_full).tolist()
            outs.append([tok[r, :n] for r, n in enumerate(sel)])       # views of this call's own fresh CPU tensor
        return outs

    def forward(self, src_inputs=None, src_sizes=None, first_tokens=None, src_mask=None, src_langs=None, tgt_langs=None,
                pad_idx=None, max_len: int = None, unpad_output: bool = True, beam_width: int = None, images=None,
                proposals=None, image_embed=None):
        # the model's device becomes the thread's current device for the call (function attributes, graph capture and plain
        # launches all act on the CURRENT device; DataParallel's worker threads set it, a plain caller may not have)
        with torch.cuda.device(self._model_device()):
            g = self._prepare(src_inputs, src_sizes, first_tokens, src_mask, src_langs, tgt_langs, pad_idx, max_len, unpad_output,
                              beam_width, images, proposals, image_embed)
            return self._decode_groups([g])[0]

    def _model_device(self):
        model = self.seq2seq_model.module if hasattr(self.seq2seq_model, "module") else self.seq2seq_model
        device = model.encoder.embeddings.word_embeddings.weight.device
        if device.type != "cuda":
            raise RuntimeError("imagetranslate_b200.BeamDecoder needs the model on a CUDA device (no CPU fallback)")
        return device

    # ------------------------------------------------------------------ several batches in flight
    @staticmethod
    def _unwrap(v):
        return v[0] if isinstance(v, list) else v

    def _pack(self, batches, rows):
        """Consecutive batches that can share a super-batch (same target-language decoder, beam, pad id, memory kind,
        similar memory length), up to `rows` beam rows.  Returns a list of index lists."""
        packs, cur, cur_key, cur_rows, s_lo, s_hi = [], [], None, 0, 0, 0
        for i, kw in enumerate(batches):
            first, tl = self._unwrap(kw.get("first_tokens")), self._unwrap(kw.get("tgt_langs"))
            beam = kw.get("beam_width") or self.beam_width
            src = self._unwrap(kw.get("src_inputs"))
            S = int(src.size(1)) if src is not None else 49
            prop = self._unwrap(kw.get("proposals"))
            key = (int(tl[0]), beam, kw.get("pad_idx"), src is None, kw.get("max_len"), kw.get("images") is not None,
                   None if prop is None else int(prop.shape[1]))      # PAD proposals are attended (seq2seq.py:131): no padding
            n = int(first.numel()) * beam
            ok = cur and key == cur_key and cur_rows + n <= rows and max(s_hi, S) <= self.pack_ratio * min(s_lo, S)
            if not ok:
                if cur:
                    packs.append(cur)
                cur, cur_key, cur_rows, s_lo, s_hi = [], key, 0, S, S
            cur.append(i)
            cur_rows += n
            s_lo, s_hi = min(s_lo, S), max(s_hi, S)
        if cur:
            packs.append(cur)
        return packs

    def forward_many(self, batches, lanes: int = 3, rows: int = 4096):
        """Decodes a list of batches (each a dict of `forward` keyword arguments) and returns their outputs in order.

        Every batch is decoded exactly as `forward` would decode it (same loop bound, termination rule and memory per
        reference batch).  In the fp32 parity mode the results are invariant to `lanes` and `rows`; in the 16-bit modes they are
        deterministic for a fixed `rows` and independent of `lanes`, but NOT invariant to `rows`: the GEMM tile height,
        split-K factor and attention staging are chosen by launch size, so the summation order -- and with it the winner of a
        near-tie -- can change with the super-batch size.  Otherwise:
          * consecutive compatible batches are decoded in lockstep as one SUPER-BATCH of up to `rows` beam rows: one
            kernel chain and one pass over the 191 MB of decoder + vocabulary weights per step serves all of them
            (a single reference batch at --batch 4000 has 120-940 rows and leaves the step latency-bound);
          * up to `lanes` super-batches are in flight on separate CUDA streams, each with its own workspace / KV cache /
            captured step graph, which fills the SMs the dependent-kernel chain of one super-batch leaves idle.
        This replaces the reference's thread-per-GPU `DataParallelModel` fan-out (src/parallel.py:70-148) on ONE device."""
        batches = list(batches)
        if not batches:
            return []
        model = self.seq2seq_model.module if hasattr(self.seq2seq_model, "module") else self.seq2seq_model
        wemb = model.encoder.embeddings.word_embeddings.weight
        device = wemb.device
        if device.type != "cuda":
            raise RuntimeError("imagetranslate_b200.BeamDecoder needs the model on a CUDA device (no CPU fallback)")
        with torch.cuda.device(device):
            return self._forward_many(batches, lanes, rows, model, wemb, device)

    def _forward_many(self, batches, lanes, rows, model, wemb, device):
        packs = self._pack(batches, max(int(rows), 1))
        results = [None] * len(batches)

        def run_pack(idxs):
            groups = [self._prepare(**batches[i]) for i in idxs]
            for i, o in zip(idxs, self._decode_groups(groups)):
                results[i] = o

        lanes = max(1, min(int(lanes), len(packs)))
        if lanes > 1:
            # longest super-batch first: the lanes take super-batches from one queue, so the last one to start should be
            # the shortest, or one lane ends up decoding the most expensive super-batch alone (results are stored by batch
            # index and do not depend on the order)
            from .parallel import batch_cost

            def cost(idxs):
                c = 0.0
                for i in idxs:
                    kw = batches[i]
                    src = self._unwrap(kw.get("src_inputs"))
                    n = int(self._unwrap(kw["first_tokens"]).numel())
                    S = int(src.size(1)) if src is not None else 49
                    c += batch_cost(n, S, kw.get("beam_width") or self.beam_width, self.max_len_a, self.max_len_b) if kw.get("max_len") is None \
                        else float(kw["max_len"]) * n
                return c

            packs = sorted(packs, key=lambda idxs: (-cost(idxs), idxs[0]))
        if lanes == 1:
            for idxs in packs:
                run_pack(idxs)
            return results
        if not hasattr(self, "_streams") or len(self._streams) < lanes:
            self._streams = [torch.cuda.Stream(device=device) for _ in range(lanes)]
            self._pool = ThreadPoolExecutor(max_workers=lanes)
        code = self._dtype_code(wemb)
        for kw in batches:                     # pack the weights once, on the calling thread, before the lanes start
            self._weights(model, int(self._unwrap(kw["tgt_langs"])[0]), code, device)
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(device))
        nxt = [0]

        def lane_main(idx):
            self._lane.idx = idx
            stream = self._streams[idx]
            with torch.cuda.device(device), torch.cuda.stream(stream):
                stream.wait_event(ready)
                while True:
                    with self._lock:
                        i = nxt[0]
                        nxt[0] += 1
                    if i >= len(packs):
                        break
                    run_pack(packs[i])
                done = torch.cuda.Event()
                done.record(stream)
            return done

        futs = [self._pool.submit(lane_main, i) for i in range(lanes)]
        err = None
        for f in futs:
            try:
                torch.cuda.current_stream(device).wait_event(f.result())
            except Exception as e:             # a lane's RuntimeError surfaces after all lanes stopped
                err = err or e
        if err is not None:
            raise err
        return results
