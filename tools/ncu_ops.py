"""A few eager launches of single ops at a chosen row count, for `ncu --set full` captures (not a test, not the bench).
Usage: python tools/ncu_ops.py [B] [S] [t] [ops]   ops: comma list of self,cross,qkv,o,ffn1,ffn2,vocab,addln"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from imagetranslate_b200 import _lib  # noqa: E402

DEV = "cuda:0"
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
S = int(sys.argv[2]) if len(sys.argv) > 2 else 40
t = int(sys.argv[3]) if len(sys.argv) > 3 else 30
OPS = (sys.argv[4] if len(sys.argv) > 4 else "self,cross,ffn1,ffn2,o").split(",")
DT, ACT = _lib.BF16, torch.bfloat16
beam, d, H, F, L, V, T = 4, 768, 12, 3072, 6, 60000, int(1.3 * S + 5)
M = B * beam
lib = _lib.lib()
REP = 3


def st():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def rnd(*shape, dt=torch.float32, scale=1.0):
    return (torch.randn(*shape, device=DEV) * scale).to(dt)


ctl = torch.tensor([t, 0, t - 1, 0, 0, 0, 1000, 0], dtype=torch.int32, device=DEV)


def gemm(N, K, epi):
    W, bias, A = rnd(N, K, dt=ACT, scale=0.02), rnd(N), rnd(M, K, dt=ACT)
    out = torch.empty(M, (N + 3) // 4 * 4, dtype=torch.float32 if epi == _lib.EPI_BIAS_F32 else ACT, device=DEV)
    for _ in range(REP):
        _lib.check(lib.itb_linear(DT, _lib.GEMM_AUTO, epi, _lib.ptr(A), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(out), M, N, K,
                                  out.shape[1], st()))


if "qkv" in OPS:
    gemm(3 * d, d, _lib.EPI_BIAS_F32)
if "o" in OPS:
    gemm(d, d, _lib.EPI_BIAS_F32)
if "ffn1" in OPS:
    gemm(F, d, _lib.EPI_BIAS_GELU_ACT)
if "ffn2" in OPS:
    gemm(d, F, _lib.EPI_BIAS_F32)
if "vocab" in OPS:
    gemm(V, d, _lib.EPI_BIAS_F32)
if "vocabf" in OPS:      # fused vocabulary projection (logits never leave the SM)
    W, bias, A = rnd(V, d, dt=ACT, scale=0.02), rnd(V), rnd(M, d, dt=ACT)
    nc, KT = lib.itb_vocab_chunks(V), lib.itb_topk_capacity(beam)
    cv, ci = torch.empty(M, nc, KT, device=DEV), torch.empty(M, nc, KT, dtype=torch.int32, device=DEV)
    cm, cs = torch.empty(M, nc, device=DEV), torch.empty(M, nc, device=DEV)
    for _ in range(REP):
        _lib.check(lib.itb_vocab_partials(DT, beam, _lib.ptr(A), _lib.ptr(W), _lib.ptr(bias), M, V, d, _lib.ptr(cv), _lib.ptr(ci),
                                          _lib.ptr(cm), _lib.ptr(cs), None, st()))
if "addln" in OPS:
    y, xf, xa, g_, b_ = rnd(M, d), rnd(M, d), rnd(M, d, dt=ACT), rnd(d), rnd(d)
    for _ in range(REP):
        _lib.check(lib.itb_add_ln(DT, _lib.ptr(ctl), _lib.ptr(y), 1, _lib.ptr(g_), _lib.ptr(b_), 1e-12, _lib.ptr(xf), _lib.ptr(xa), M, d, st()))
qkv = rnd(M, 3 * d)
ctx = torch.empty(M, d, dtype=ACT, device=DEV)
if "self" in OPS:
    kv = rnd(M, T, 2, d, dt=ACT)
    anc = torch.randint(0, beam, (2, M, T), device=DEV)
    SHARE = max(t - 6, 0)
    anc[:, :, :SHARE] = anc.view(2, B, beam, T)[:, :, :1, :SHARE].expand(2, B, beam, SHARE).reshape(2, M, SHARE)
    anc = anc.to(torch.uint8)
    for _ in range(REP):
        _lib.check(lib.itb_self_attention(DT, _lib.ptr(ctl), None, _lib.ptr(qkv), _lib.ptr(kv), _lib.ptr(anc), _lib.ptr(ctx), M, beam, d, H, T, st()))
if "cross" in OPS:
    ckv = rnd(B, S, L, 2, d, dt=ACT)
    mask = torch.ones(B, S, dtype=torch.uint8, device=DEV)
    for _ in range(REP):
        _lib.check(lib.itb_cross_attention(DT, _lib.ptr(ctl), None, None, _lib.ptr(qkv), 3 * d, _lib.ptr(ckv), 2, _lib.ptr(mask), _lib.ptr(ctx),
                                           M, beam, d, H, S, L, st()))
torch.cuda.synchronize()
print("ok")
