"""Per-kernel timing of the decode step on the GPU box (not a test, not the bench): every op of one search step is
captured REP times back-to-back into a CUDA graph and replayed, which gives the in-graph cost per launch (kernel body +
dependent-launch gap), the number that adds up to the step time.  Usage: python tools/kernel_bench.py [B] [S] [t] [dtype]"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from imagetranslate_b200 import _lib  # noqa: E402

DEV = "cuda:0"
B = int(sys.argv[1]) if len(sys.argv) > 1 else 61
S = int(sys.argv[2]) if len(sys.argv) > 2 else 64
t = int(sys.argv[3]) if len(sys.argv) > 3 else 44
DT = {"bf16": _lib.BF16, "fp16": _lib.F16, "fp32": _lib.F32}[sys.argv[4] if len(sys.argv) > 4 else "bf16"]
ACT = {_lib.F32: torch.float32, _lib.BF16: torch.bfloat16, _lib.F16: torch.float16}[DT]
beam, d, H, F, L, V, T = 4, 768, 12, 3072, 6, 60000, int(1.3 * S + 5)
M = B * beam
REP = 24
lib = _lib.lib()


def bench(name, fn, bytes_alg=0, nbuf=1):
    """fn(i) enqueues launch i on the current stream (i selects one of nbuf rotating operand sets)."""
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(3):
            fn(i % nbuf)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for i in range(REP):
                fn(i % nbuf)
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(10):
            g.replay()
        e1.record()
        e1.synchronize()
    us = e0.elapsed_time(e1) * 1000.0 / (10 * REP)
    gbs = bytes_alg / us / 1e3 if bytes_alg else 0.0
    print("%-34s %8.2f us/launch   %8.1f GB/s (algorithmic)" % (name, us, gbs))
    return us


def st():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def rnd(*shape, dt=torch.float32, scale=1.0):
    return (torch.randn(*shape, device=DEV) * scale).to(dt)


ctl = torch.tensor([t, 0, t - 1, 0, 0, 0, 1000, 0], dtype=torch.int32, device=DEV)
total = 0.0
NB = 6      # rotate over 6 weight copies (one per layer) so the weights do not sit in L2 more than they would in a real step


def gemm(name, N, K, epi, count):
    global total
    W = [rnd(N, K, dt=ACT, scale=0.02) for _ in range(NB if N < 10000 else 1)]
    bias = rnd(N)
    A = rnd(M, K, dt=ACT)
    out = torch.empty(M, (N + 3) // 4 * 4, dtype=torch.float32 if epi == _lib.EPI_BIAS_F32 else ACT, device=DEV)
    ldo = out.shape[1]
    us = bench("%s M=%d N=%d K=%d" % (name, M, N, K),
               lambda i: _lib.check(lib.itb_linear(DT, _lib.GEMM_AUTO, epi, _lib.ptr(A), _lib.ptr(W[i % len(W)]), _lib.ptr(bias),
                                                   _lib.ptr(out), M, N, K, ldo, st())),
               bytes_alg=N * K * W[0].element_size() + M * K * A.element_size() + M * N * out.element_size(), nbuf=len(W))
    total += us * count


gemm("qkv", 3 * d, d, _lib.EPI_BIAS_F32, L)
gemm("o/cq/co", d, d, _lib.EPI_BIAS_F32, 3 * L)
gemm("ffn1+gelu", F, d, _lib.EPI_BIAS_GELU_ACT, L)
gemm("ffn2", d, F, _lib.EPI_BIAS_F32, L)
gemm("vocab", V, d, _lib.EPI_BIAS_F32, 1)

# residual + LayerNorm
y, xf, xa, g_, b_ = rnd(M, d), rnd(M, d), rnd(M, d, dt=ACT), rnd(d), rnd(d)
total += 3 * L * bench("add_ln", lambda i: _lib.check(lib.itb_add_ln(DT, _lib.ptr(ctl), _lib.ptr(y), 1, _lib.ptr(g_), _lib.ptr(b_), 1e-12,
                                                                      _lib.ptr(xf), _lib.ptr(xa), M, d, st())), bytes_alg=M * d * 14)
# self attention at position t-1
qkv = rnd(M, 3 * d)
kv = [rnd(M, T, 2, d, dt=ACT) for _ in range(NB)]
anc = torch.randint(0, beam, (2, M, T), device=DEV)
SHARE = int(os.environ.get("KB_SHARE", max(t - 6, 0)))      # positions before which the slots of a sentence share rows
anc[:, :, :SHARE] = anc.view(2, B, beam, T)[:, :, :1, :SHARE].expand(2, B, beam, SHARE).reshape(2, M, SHARE)
anc = anc.to(torch.uint8)
ctx = torch.empty(M, d, dtype=ACT, device=DEV)
esz = kv[0].element_size()
total += L * bench("self_attn t=%d" % t, lambda i: _lib.check(lib.itb_self_attention(DT, _lib.ptr(ctl), None, _lib.ptr(qkv), _lib.ptr(kv[i]),
                                                                                      _lib.ptr(anc), _lib.ptr(ctx), M, beam, d, H, T, st())),
                   bytes_alg=M * t * 2 * d * esz, nbuf=NB)
# cross attention
ckv = rnd(B, S, L, 2, d, dt=ACT)
mask = torch.ones(B, S, dtype=torch.uint8, device=DEV)
total += L * bench("cross_attn S=%d" % S, lambda i: _lib.check(lib.itb_cross_attention(DT, _lib.ptr(ctl), None, None, _lib.ptr(qkv), 3 * d, _lib.ptr(ckv),
                                                                                        i % L, _lib.ptr(mask), _lib.ptr(ctx), M, beam, d, H,
                                                                                        S, L, st())),
                   bytes_alg=B * S * 2 * d * esz, nbuf=NB)
# the same bytes with one layer per buffer (rows 3 KB apart instead of L x 3 KB): what a layer-major cross-KV layout would give
ckv1 = [rnd(B, S, 1, 2, d, dt=ACT) for _ in range(L)]
bench("cross_attn S=%d, layer-major rows" % S, lambda i: _lib.check(lib.itb_cross_attention(DT, _lib.ptr(ctl), None, None, _lib.ptr(qkv), 3 * d,
                                                                                             _lib.ptr(ckv1[i % L]), 0, _lib.ptr(mask), _lib.ptr(ctx),
                                                                                             M, beam, d, H, S, 1, st())),
      bytes_alg=B * S * 2 * d * esz, nbuf=NB)
print("sum over one search step (6 layers + vocab GEMM, without embed / beam kernel): %.1f us" % total)
