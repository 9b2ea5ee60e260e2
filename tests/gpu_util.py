"""Helpers for the -m gpu parity tests: thin callers of the C ABI plus torch/oracle references."""
import ctypes as C

import torch

from imagetranslate_b200 import _lib

DEV = "cuda:0"
ACT = {_lib.F32: torch.float32, _lib.BF16: torch.bfloat16, _lib.F16: torch.float16}


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def make_ctl(step=1, done=0, cur=0, max_len=1000):
    return torch.tensor([step, done, step - 1, cur, 0, 0, max_len, 0], dtype=torch.int32, device=DEV)


def linear(dtype, impl, epi, A, W, bias):
    M, K = A.shape
    N = W.shape[0]
    odt = torch.float32 if epi == _lib.EPI_BIAS_F32 else ACT[dtype]
    out = torch.full((M, N), float("nan"), dtype=torch.float32, device=DEV).to(odt)
    _lib.check(_lib.lib().itb_linear(dtype, impl, epi, _lib.ptr(A), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(out), M, N, K, N,
                                     stream()), "linear")
    torch.cuda.synchronize()
    return out


def rel_err(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / (b.norm() + 1e-30))


class BeamHarness(object):
    """Owns the search state + scratch of itb_beam_step for synthetic-logit tests."""

    def __init__(self, B, beam, V, T_max, max_lens, first_tokens, max_len, pad=0, eos=4, alpha=0.8, n_chunk=4):
        l = _lib.lib()
        self.B, self.beam, self.V, self.T = B, beam, V, T_max
        M = B * beam
        self.M, self.V4 = M, (V + 3) // 4 * 4
        KT = l.itb_topk_capacity(beam)
        i32, u8, f32 = torch.int32, torch.uint8, torch.float32
        def z(n, dt):     # poisoned, not zero: kernels must never rely on fresh allocations
            if dt == torch.float32:
                return torch.full((n,), 3.0e38, dtype=dt, device=DEV)
            return torch.full((n,), 77, dtype=dt, device=DEV)
        self.t = dict(ctl=z(8, i32), tokens=z(2 * M * T_max, i32), anc=z(2 * M * T_max, u8), scores=z(2 * M, f32),
                      sizes=z(2 * M, f32), fin=z(2 * M, u8), logits=z(M * self.V4, f32), cand_val=z(M * n_chunk * KT, f32),
                      cand_idx=z(M * n_chunk * KT, i32), chunk_max=z(M * n_chunk, f32), chunk_sum=z(M * n_chunk, f32),
                      sent_counter=z(B, i32), max_lens=torch.tensor(max_lens, dtype=i32, device=DEV),
                      sent_grp=torch.zeros(B, dtype=i32, device=DEV), grp=z(8, i32), sent_done=z(B, i32))
        c = _lib.Config()
        c.dtype, c.B, c.beam, c.V, c.T_max, c.S, c.L, c.d, c.H, c.F = _lib.F32, B, beam, V, T_max, 1, 1, 64, 1, 64
        c.pad_id, c.eos_id, c.n_chunk, c.len_alpha, c.G = pad, eos, n_chunk, alpha, 1
        for k, v in self.t.items():
            setattr(c, k, v.data_ptr())
        self.cfg = c
        first = torch.tensor(first_tokens, dtype=i32, device=DEV)
        self.grp_max_len = torch.tensor([max_len], dtype=i32, device=DEV)
        _lib.check(l.itb_init_state(C.byref(c), _lib.ptr(first), _lib.ptr(self.grp_max_len), stream()), "init_state")
        torch.cuda.synchronize()

    def ctl(self):
        return self.t["ctl"].cpu().tolist()

    def step(self, logits):
        self.t["logits"].view(self.M, self.V4)[:, :self.V].copy_(logits.to(DEV))
        _lib.check(_lib.lib().itb_beam_step(C.byref(self.cfg), stream()), "beam_step")
        torch.cuda.synchronize()

    def step_fused(self, logits):
        """The 16-bit modes' flow: the per-chunk partials that the vocabulary GEMM's epilogue would emit (chunks of 128
        entries: maximum, sum exp(x - max), top-KT by value desc / index asc) are computed here with torch from `logits`,
        then itb_beam_step runs with fused_vocab = 1 (phase 2 only, one CTA per sentence)."""
        l = _lib.lib()
        nc, KT = l.itb_vocab_chunks(self.V), l.itb_topk_capacity(self.beam)
        assert nc == self.cfg.n_chunk
        x = torch.full((self.M, nc * 128), float("-inf"), device=DEV)
        x[:, :self.V] = logits.to(DEV)
        x = x.view(self.M, nc, 128)
        cm = x.max(dim=2).values
        safe = torch.where(torch.isinf(cm), torch.zeros_like(cm), cm)
        cs = torch.exp(x - safe[:, :, None]).sum(dim=2)
        order = torch.argsort(x, dim=2, descending=True, stable=True)[:, :, :KT]
        cv = torch.gather(x, 2, order)
        ci = (order + (torch.arange(nc, device=DEV) * 128)[None, :, None]).to(torch.int32)
        ci = torch.where(torch.isinf(cv), torch.full_like(ci, -1), ci)
        self.t["chunk_max"].copy_(cm.reshape(-1))
        self.t["chunk_sum"].copy_(cs.reshape(-1))
        self.t["cand_val"].copy_(cv.reshape(-1))
        self.t["cand_idx"].copy_(ci.reshape(-1))
        self.cfg.fused_vocab = 1
        _lib.check(l.itb_beam_step(C.byref(self.cfg), stream()), "beam_step")
        torch.cuda.synchronize()

    def state(self):
        ctl = self.ctl()
        cur, n = ctl[3], ctl[2] + 1
        tok = self.t["tokens"].view(2, self.M, self.T)[cur, :, :n].cpu().long().view(self.B, self.beam, n)
        sc = self.t["scores"].view(2, self.M)[cur].cpu().view(self.B, self.beam)
        sz = self.t["sizes"].view(2, self.M)[cur].cpu().view(self.B, self.beam)
        anc = self.t["anc"].view(2, self.M, self.T)[cur, :, :max(n - 1, 0)].cpu().long().view(self.B, self.beam, -1)
        return tok, sc, sz, anc
