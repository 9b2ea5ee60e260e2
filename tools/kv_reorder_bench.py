"""KV-cache reorder variants of SURVEY.md §7.2-6 / BASELINE.json configs[3] (sources of 256 tokens, beam 1/4/8/12), measured
side by side on the GPU box (not a test, not the bench):

  indirection  what the decode path ships: the cache is never moved; the beam kernel maintains an ancestor table and the
               self-attention kernel reads K,V through it, loading the prefix the beam slots share ONCE per sentence.
  physical     what a literal KV-cached port of reference src/seq_gen.py:217-219 does: gather every row's whole cache by
               back-pointer each step (itb_kv_gather over all layers), after which every row owns a private copy and attention
               reads M full rows (identity ancestor table, nothing shared).

Each variant = the launches one search step needs for ALL layers, captured REP times into a CUDA graph and replayed.
Usage: python tools/kv_reorder_bench.py > profiles/r2_kv_reorder_variants.txt
"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from imagetranslate_b200 import _lib  # noqa: E402

DEV = "cuda:0"
lib = _lib.lib()
d, H, L, S = 768, 12, 6, 256
T = 352
REP = 8


def st():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def timed(fn):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(REP):
                fn()
        for _ in range(2):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(5):
            g.replay()
        e1.record()
        e1.synchronize()
    return e0.elapsed_time(e1) * 1000.0 / (5 * REP)


print("%-6s %-5s %-5s %-6s | %-30s | %-44s | %s" % ("B", "beam", "rows", "t", "indirection: attention (us/step)", "physical: gather + attention = total (us/step)", "ratio"))
for B in (15, 64):
    for beam in (1, 4, 8):
        M = B * beam
        for t in (170, 336):
            dt = _lib.F16
            ctl = torch.tensor([t, 0, t - 1, 0, 0, 0, 1000, 0], dtype=torch.int32, device=DEV)
            qkv = torch.randn(M, 3 * d, device=DEV)
            kv = torch.randn(L, M, T, 2, d, device=DEV).to(torch.float16)
            kv2 = torch.empty_like(kv)
            torch.cuda.synchronize()
            ctx = torch.empty(M, d, dtype=torch.float16, device=DEV)
            # realistic ancestor table: the slots of a sentence share their lineage up to the last few positions
            anc = torch.randint(0, beam, (2, M, T), device=DEV)
            share = max(t - 6, 0)
            anc[:, :, :share] = anc.view(2, B, beam, T)[:, :, :1, :share].expand(2, B, beam, share).reshape(2, M, share)
            anc = anc.to(torch.uint8)
            ident = (torch.arange(M, device=DEV) % beam).to(torch.uint8).view(1, M, 1).expand(2, M, T).contiguous()
            src_row = (torch.arange(M, device=DEV) // beam * beam + torch.randint(0, beam, (M,), device=DEV)).to(torch.int32)
            esz = 2

            def attn(table):
                for l in range(L):
                    _lib.check(lib.itb_self_attention(dt, _lib.ptr(ctl), None, _lib.ptr(qkv), _lib.ptr(kv[l]), _lib.ptr(table), _lib.ptr(ctx),
                                                      M, beam, d, H, T, st()))

            def gather():
                _lib.check(lib.itb_kv_gather(dt, _lib.ptr(kv), _lib.ptr(kv2), _lib.ptr(src_row), M, T, t - 1, d, L, st()))

            t_ind = timed(lambda: attn(anc))
            t_gat = timed(gather)
            t_att = timed(lambda: attn(ident))
            by_ind = L * (B * share + M * (t - share)) * 2 * d * esz
            by_phy = L * M * (t - 1) * 2 * d * esz * 2 + L * M * t * 2 * d * esz
            print("%-6d %-5d %-5d %-6d | %8.1f  (%6.1f MB K,V read)   | %8.1f + %8.1f = %8.1f  (%7.1f MB moved)  | %.2fx" % (
                B, beam, M, t, t_ind, by_ind / 1e6, t_gat, t_att, t_gat + t_att, by_phy / 1e6, (t_gat + t_att) / t_ind))
