"""Drop-in `BeamDecoder` (reference src/seq_gen.py:27-242) over the B200 decode engine.

Same constructor and `forward` signature, same list-unwrapping of DataParallel-style kwargs
(:57-72), same max_len / max_lens rules (:113-121), same return value (list of 1-D CPU LongTensors
that start with the first/language token and exclude EOS; `unpad_output=False` returns full rows).
The search itself (decoder step, vocabulary projection, log-softmax, scoring, top-k, back-pointers,
termination) runs in CUDA kernels -- see include/itb.h.  The encoder memory still comes from the
model's own `encode` (:96,99,106), exactly as in the reference.
"""
import threading
from concurrent.futures import ThreadPoolExecutor

import torch
import torch.nn as nn

from . import _lib
from .engine import DecodeEngine, PackedDecoder


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

    def refresh_weights(self):
        """Drop packed weights (call after loading a new checkpoint into the wrapped model)."""
        self._packed.clear()

    # ------------------------------------------------------------------ forward
    def forward(self, src_inputs=None, src_sizes=None, first_tokens=None, src_mask=None, src_langs=None, tgt_langs=None,
                pad_idx=None, max_len: int = None, unpad_output: bool = True, beam_width: int = None, images=None,
                proposals=None, image_embed=None):
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
        if getattr(model, "use_proposals", False):
            raise RuntimeError("lexical proposals (attend_proposal) are not supported by the B200 decode path yet")
        if src_inputs is not None and images is not None:
            raise RuntimeError("dual text+image stream decoding is not supported by the B200 decode path yet")
        batch_lang = int(tgt_langs[0])

        with torch.no_grad():
            if src_inputs is not None:                            # :94-96
                src_mask = src_mask.to(device)
                langs2d = src_langs.unsqueeze(-1).expand(-1, src_inputs.size(-1))
                memory = model.encode(src_inputs, src_mask, langs2d)[0]
                mem_mask = src_mask
            else:
                if image_embed is None:                           # :97-99
                    images = images.to(device)
                    if wemb.dtype == torch.float16:
                        images = images.half()
                    memory, obj_feat = model.encode(images=images)
                    if obj_feat is not None:
                        raise RuntimeError("object-feature stream (--obj) is not supported by the B200 decode path yet")
                else:                                             # :100-103
                    memory = image_embed.to(device)
                if max_len is None and images is not None:
                    max_len = 512                                 # :90-91
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
            max_lens = torch.tensor([mlf(int(x)) for x in src_sizes], dtype=torch.int64)   # :121
        max_len = min(int(max_len), n_pos)

        code = self._dtype_code(wemb)
        eng = self._engine(device, code)
        pw = self._weights(model, batch_lang, code, device)
        tok, lens, lens_full = eng.decode(pw, memory, mem_mask, first_tokens, tgt_langs, max_lens, max_len, beam_width,
                                          int(pad_idx), int(eos), float(self.len_penalty_ratio))
        with self._lock:
            self.kernel_launches += eng.last_steps * eng.launches_per_step + 4
        tok = tok.to(torch.int64)
        sel = lens if unpad_output else lens_full
        return [tok[r, :int(sel[r])].clone() for r in range(B)]

    # ------------------------------------------------------------------ several batches in flight
    def forward_many(self, batches, lanes: int = 4):
        """Decodes a list of batches (each a dict of `forward` keyword arguments) and returns their outputs in order.

        Every batch is decoded exactly as `forward` would (whole reference batches, never re-batched, so results do
        not depend on `lanes`), but up to `lanes` batches are in flight on separate CUDA streams, each with its own
        workspace / KV cache / captured step graph.  A single search step at the reference batch sizes is a chain of
        ~70 short latency-bound kernels that leaves most SMs idle; independent batches fill them.  This replaces the
        reference's one-thread-per-GPU `DataParallelModel` fan-out (src/parallel.py:70-148) on ONE device."""
        batches = list(batches)
        if lanes <= 1 or len(batches) <= 1:
            return [self.forward(**kw) for kw in batches]
        model = self.seq2seq_model.module if hasattr(self.seq2seq_model, "module") else self.seq2seq_model
        device = model.encoder.embeddings.word_embeddings.weight.device
        if device.type != "cuda":
            raise RuntimeError("imagetranslate_b200.BeamDecoder needs the model on a CUDA device (no CPU fallback)")
        lanes = min(lanes, len(batches))
        if not hasattr(self, "_streams") or len(self._streams) < lanes:
            self._streams = [torch.cuda.Stream(device=device) for _ in range(lanes)]
            self._pool = ThreadPoolExecutor(max_workers=lanes)
        # pack the weights once, on the calling thread, before the lanes start
        wemb = model.encoder.embeddings.word_embeddings.weight
        code = self._dtype_code(wemb)
        for kw in batches:
            tl = kw["tgt_langs"][0] if isinstance(kw["tgt_langs"], list) else kw["tgt_langs"]
            self._weights(model, int(tl[0]), code, device)
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(device))
        results = [None] * len(batches)
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
                    if i >= len(batches):
                        break
                    results[i] = self.forward(**batches[i])
                done = torch.cuda.Event()
                done.record(stream)
            return done

        futs = [self._pool.submit(lane_main, i) for i in range(lanes)]
        for f in futs:
            torch.cuda.current_stream(device).wait_event(f.result())     # re-raises a lane's RuntimeError
        return results
