"""Kernel-level parity (-m gpu): every C-ABI op against a plain torch / oracle restatement of the same math."""
import ctypes as C

import pytest
import torch

from imagetranslate_b200 import _lib
from oracle import restated as R

from gpu_util import ACT, DEV, BeamHarness, linear, make_ctl, rel_err, stream

pytestmark = pytest.mark.gpu


def _rand(shape, seed, scale=1.0, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(DEV).to(dtype)


# ------------------------------------------------------------------------------------------------ GEMM
@pytest.mark.parametrize("M,N,K", [(5, 200, 64), (61, 768, 768), (244, 2304, 768), (70, 768, 3072), (130, 1000, 768)])
@pytest.mark.parametrize("epi", [_lib.EPI_BIAS_F32, _lib.EPI_BIAS_GELU_ACT, _lib.EPI_BIAS_ACT])
def test_linear_f32_cuda_cores(M, N, K, epi):
    A, W, b = _rand((M, K), 1), _rand((N, K), 2, 0.05), _rand((N,), 3)
    ref = A.double() @ W.double().t() + b.double()
    if epi == _lib.EPI_BIAS_GELU_ACT:
        ref = torch.nn.functional.gelu(ref)
    out = linear(_lib.F32, _lib.GEMM_AUTO, epi, A, W, b)
    assert rel_err(out, ref) < 2e-6


@pytest.mark.parametrize("impl", [_lib.GEMM_SIMT, _lib.GEMM_TCGEN05], ids=["simt", "tcgen05"])
@pytest.mark.parametrize("M,N,K", [(7, 256, 64), (32, 768, 768), (60, 2304, 768), (100, 768, 3072), (244, 3072, 768),
                                   (300, 1000, 768), (940, 768, 768), (244, 60000, 768), (16, 9216, 768),
                                   (4096, 768, 768), (4000, 768, 3072), (2048, 3072, 768)])
def test_linear_bf16(impl, M, N, K):
    A, W, b = _rand((M, K), 4, 1.0, torch.bfloat16), _rand((N, K), 5, 0.05, torch.bfloat16), _rand((N,), 6)
    ref = A.double() @ W.double().t() + b.double()
    out = linear(_lib.BF16, impl, _lib.EPI_BIAS_F32, A, W, b)
    assert torch.isfinite(out).all()
    assert rel_err(out, ref) < 2e-5          # bf16 products are exact in fp32; only the summation order differs
    out2 = linear(_lib.BF16, impl, _lib.EPI_BIAS_GELU_ACT, A, W, b)
    assert rel_err(out2.float(), torch.nn.functional.gelu(ref)) < 6e-3
    out3 = linear(_lib.BF16, impl, _lib.EPI_BIAS_ACT, A, W, b)
    assert rel_err(out3.float(), ref) < 6e-3


@pytest.mark.parametrize("impl", [_lib.GEMM_SIMT, _lib.GEMM_TCGEN05], ids=["simt", "tcgen05"])
@pytest.mark.parametrize("M,N,K", [(32, 768, 768), (244, 3072, 768), (100, 768, 3072), (300, 1000, 768)])
def test_linear_fp16(impl, M, N, K):
    """ITB_F16: fp16 weights and activations (UMMA kind::f16 with both operand formats f16)."""
    A, W, b = _rand((M, K), 4, 1.0, torch.float16), _rand((N, K), 5, 0.05, torch.float16), _rand((N,), 6)
    ref = A.double() @ W.double().t() + b.double()
    out = linear(_lib.F16, impl, _lib.EPI_BIAS_F32, A, W, b)
    assert torch.isfinite(out).all()
    assert rel_err(out, ref) < 2e-5
    out2 = linear(_lib.F16, impl, _lib.EPI_BIAS_GELU_ACT, A, W, b)
    assert out2.dtype == torch.float16 and rel_err(out2.float(), torch.nn.functional.gelu(ref)) < 1e-3


@pytest.mark.parametrize("dtype", [_lib.BF16, _lib.F16], ids=["bf16", "fp16"])
@pytest.mark.parametrize("rows,V,beam", [(5, 300, 4), (130, 1000, 3), (244, 60000, 4), (40, 5003, 8), (33, 777, 12), (260, 4096, 1),
                                         (700, 3000, 5), (520, 1500, 4), (1024, 2200, 4),
                                         (4096, 60000, 4)])      # last: the bench's full size (BASELINE configs[1])
def test_fused_vocab_projection_partials(dtype, rows, V, beam):
    """itb_vocab_partials: per-chunk max / sum-exp / top-KT emitted by the GEMM epilogue reproduce log_softmax's
    log-sum-exp and the row top-k (value desc, index asc) of the full logits; the optional logits output equals the GEMM."""
    lib, K = _lib.lib(), 768
    act = ACT[dtype]
    x, w, b = _rand((rows, K), 31, 1.0, act), _rand((V, K), 32, 0.05, act), _rand((V,), 33)
    nc, KT, V4 = lib.itb_vocab_chunks(V), lib.itb_topk_capacity(beam), (V + 3) // 4 * 4
    assert nc == 2 * ((V + 255) // 256)
    cand_val = torch.full((rows, nc, KT), float("nan"), device=DEV)
    cand_idx = torch.full((rows, nc, KT), -7, dtype=torch.int32, device=DEV)
    cmax, csum = torch.full((rows, nc), float("nan"), device=DEV), torch.full((rows, nc), float("nan"), device=DEV)
    logits = torch.full((rows, V4), float("nan"), device=DEV)
    for lg in (logits, None):
        _lib.check(lib.itb_vocab_partials(dtype, beam, _lib.ptr(x), _lib.ptr(w), _lib.ptr(b), rows, V, K, _lib.ptr(cand_val),
                                          _lib.ptr(cand_idx), _lib.ptr(cmax), _lib.ptr(csum), _lib.ptr(lg), stream()), "vocab_partials")
        torch.cuda.synchronize()
    ref = (x.double() @ w.double().t() + b.double())
    got = logits[:, :V]
    assert torch.isfinite(got).all() and rel_err(got, ref) < 2e-5
    # log-sum-exp from the partials (empty chunks carry max = -inf, sum = 0)
    mm = cmax.max(dim=1, keepdim=True).values
    lse = mm.squeeze(1).double() + torch.log((csum.double() * torch.exp((cmax - mm).double())).sum(1))
    assert float((lse - torch.logsumexp(got.double(), 1)).abs().max()) < 2e-5
    # chunk maxima are the maxima of their 128-entry slices of the kernel's own logits
    pad = torch.full((rows, nc * 128 - V), float("-inf"), device=DEV)
    sl = torch.cat([got, pad], 1).view(rows, nc, 128)
    assert torch.equal(cmax, sl.max(dim=2).values)
    # per-chunk candidates: sorted top-KT of the slice, lowest index first among equals, -inf / -1 when the slice runs out
    order = torch.argsort(sl, dim=2, descending=True, stable=True)[:, :, :KT]
    want_val = torch.gather(sl, 2, order)
    want_idx = (order + (torch.arange(nc, device=DEV) * 128)[None, :, None]).to(torch.int32)
    want_idx = torch.where(torch.isinf(want_val), torch.full_like(want_idx, -1), want_idx)
    assert torch.equal(cand_val, want_val)
    assert torch.equal(cand_idx, want_idx)
    # and therefore the row top-k of the candidates is the row top-k of the logits
    k = min(beam, V)
    top = torch.topk(got, k, dim=1)
    flat_v, flat_i = cand_val.view(rows, -1), cand_idx.view(rows, -1)
    tv, tp = torch.topk(flat_v, k, dim=1)
    assert torch.equal(tv, top.values)


def test_fused_vocab_projection_equal_logits_take_the_lowest_indices():
    """Rows whose logits are all equal (zero activations, zero bias) overflow the epilogue's candidate ring in some lanes
    only: the warp must fall back to the full walk together, and ties must resolve to the lowest indices."""
    lib, K, rows, V, beam = _lib.lib(), 768, 200, 1000, 4
    x, w = _rand((rows, K), 41, 1.0, torch.bfloat16), _rand((V, K), 42, 0.05, torch.bfloat16)
    x[5] = 0
    x[77] = 0
    x[131] = 0
    b = torch.zeros(V, device=DEV)
    nc, KT = lib.itb_vocab_chunks(V), lib.itb_topk_capacity(beam)
    cand_val = torch.full((rows, nc, KT), float("nan"), device=DEV)
    cand_idx = torch.full((rows, nc, KT), -7, dtype=torch.int32, device=DEV)
    cmax, csum = torch.full((rows, nc), float("nan"), device=DEV), torch.full((rows, nc), float("nan"), device=DEV)
    logits = torch.full((rows, (V + 3) // 4 * 4), float("nan"), device=DEV)
    _lib.check(lib.itb_vocab_partials(_lib.BF16, beam, _lib.ptr(x), _lib.ptr(w), _lib.ptr(b), rows, V, K, _lib.ptr(cand_val),
                                      _lib.ptr(cand_idx), _lib.ptr(cmax), _lib.ptr(csum), _lib.ptr(logits), stream()), "vocab_partials")
    torch.cuda.synchronize()
    got = logits[:, :V]
    pad = torch.full((rows, nc * 128 - V), float("-inf"), device=DEV)
    sl = torch.cat([got, pad], 1).view(rows, nc, 128)
    order = torch.argsort(sl, dim=2, descending=True, stable=True)[:, :, :KT]
    want_val = torch.gather(sl, 2, order)
    want_idx = (order + (torch.arange(nc, device=DEV) * 128)[None, :, None]).to(torch.int32)
    want_idx = torch.where(torch.isinf(want_val), torch.full_like(want_idx, -1), want_idx)
    assert torch.equal(cand_val, want_val) and torch.equal(cand_idx, want_idx)
    for r in (5, 77, 131):
        assert torch.equal(got[r], torch.zeros(V, device=DEV))
        assert cand_idx[r, 0].tolist() == [0, 1, 2, 3] and cand_idx[r, 3].tolist() == [384, 385, 386, 387]
        assert torch.equal(csum[r, :7], torch.full((7,), 128.0, device=DEV))


def test_linear_tcgen05_back_to_back_is_deterministic():
    A, W, b = _rand((244, 768), 7, 1.0, torch.bfloat16), _rand((3072, 768), 8, 0.05, torch.bfloat16), _rand((3072,), 9)
    o1 = linear(_lib.BF16, _lib.GEMM_TCGEN05, _lib.EPI_BIAS_F32, A, W, b)
    for _ in range(3):
        assert torch.equal(o1, linear(_lib.BF16, _lib.GEMM_TCGEN05, _lib.EPI_BIAS_F32, A, W, b))


# ------------------------------------------------------------------------------------------------ LN kernels
@pytest.mark.parametrize("dtype", [_lib.F32, _lib.BF16, _lib.F16])
def test_embed_ln_and_add_ln(dtype):
    M, beam, d, T, V = 12, 4, 768, 9, 50
    step = 3
    tokens = torch.randint(0, V, (2, M, T), dtype=torch.int32, device=DEV)
    lang = torch.tensor([0, 1, 1], dtype=torch.int32, device=DEV)
    word, pos, typ = _rand((V, d), 1, 0.02), _rand((512, d), 2, 0.02), _rand((2, d), 3, 0.02)
    g, b = _rand((d,), 4) * 0.1 + 1.0, _rand((d,), 5, 0.1)
    ctl = make_ctl(step=step, cur=1)
    xf = torch.zeros(M, d, device=DEV)
    xa = torch.zeros(M, d, device=DEV, dtype=ACT[dtype])
    l = _lib.lib()
    _lib.check(l.itb_embed_ln(dtype, _lib.ptr(ctl), _lib.ptr(tokens), _lib.ptr(lang), _lib.ptr(word), _lib.ptr(pos),
                              _lib.ptr(typ), _lib.ptr(g), _lib.ptr(b), 1e-12, _lib.ptr(xf), _lib.ptr(xa), M, beam, d, T,
                              stream()))
    torch.cuda.synchronize()
    tok = tokens[1, :, step - 1].long()
    e = word[tok] + pos[step - 1][None] + typ[lang.long().repeat_interleave(beam)]
    ref = torch.nn.functional.layer_norm(e, (d,), g, b, 1e-12)
    assert (xf - ref).abs().max() < 2e-5
    assert (xa.float() - ref).abs().max() < (2e-5 if dtype == _lib.F32 else 3e-2)
    y = _rand((3, M, d), 6)          # three split-K partials of the producing GEMM
    ref2 = torch.nn.functional.layer_norm(y.sum(0) + xf, (d,), g, b, 1e-12)
    _lib.check(l.itb_add_ln(dtype, _lib.ptr(ctl), _lib.ptr(y), 3, _lib.ptr(g), _lib.ptr(b), 1e-12, _lib.ptr(xf), _lib.ptr(xa),
                            M, d, stream()))
    torch.cuda.synchronize()
    assert (xf - ref2).abs().max() < 2e-5


# ------------------------------------------------------------------------------------------------ attention
@pytest.mark.parametrize("dtype", [_lib.F32, _lib.BF16, _lib.F16])
@pytest.mark.parametrize("step,T,beam", [(1, 48, 4), (2, 48, 4), (40, 48, 4), (9, 12, 4), (120, 130, 4), (70, 80, 5),
                                         (100, 112, 8), (170, 352, 12), (300, 320, 12), (33, 48, 1)])
def test_self_attention_through_ancestor_table(dtype, step, T, beam):
    """T = 12: ancestor table shorter than a warp; step 120: up to 480 virtual positions (many copy stages, table reads
    past the 32 positions held in registers); beam 5: a second group of one slot per sentence; beam 8 / 12: two / three full
    groups per sentence at long contexts (BASELINE configs[3] shapes); beam 1: greedy."""
    B, d, H = 3, 768, 12
    M, pc = B * beam, step - 1
    act = ACT[dtype]
    qkv = _rand((M, 3 * d), 11)
    kv = _rand((M, T, 2, d), 12).to(act)
    g = torch.Generator().manual_seed(13)
    anc = torch.randint(0, beam, (2, M, T), generator=g)
    # lineages of a sentence's slots share their physical rows before they split (sentence 0: at 25, sentence 1: never
    # shared, sentence 2: shared up to the current position)
    anc[:, 0:beam, :25] = anc[:, 0:1, :25]
    anc[:, 2 * beam:3 * beam, :] = anc[:, 2 * beam:2 * beam + 1, :]
    anc = anc.to(torch.uint8).to(DEV)
    ctl = make_ctl(step=step, cur=1)
    ctx = torch.zeros(M, d, device=DEV, dtype=act)
    kv_before = kv.clone()
    _lib.check(_lib.lib().itb_self_attention(dtype, _lib.ptr(ctl), None, _lib.ptr(qkv), _lib.ptr(kv), _lib.ptr(anc), _lib.ptr(ctx),
                                             M, beam, d, H, T, stream()))
    torch.cuda.synchronize()
    q, kn, vn = qkv[:, :d], qkv[:, d:2 * d].to(act), qkv[:, 2 * d:].to(act)
    # append check
    assert torch.equal(kv[:, pc, 0], kn) and torch.equal(kv[:, pc, 1], vn)
    other = torch.ones(T, dtype=torch.bool); other[pc] = False
    assert torch.equal(kv[:, other], kv_before[:, other])
    rows = torch.arange(M, device=DEV)
    base = (rows // beam) * beam
    ref = torch.zeros(M, d, device=DEV, dtype=torch.float64)
    for r in range(M):
        phys = base[r] + anc[1, r, :pc].long()
        K = torch.cat([kv_before[phys, torch.arange(pc), 0], kn[r:r + 1]], 0).double().view(pc + 1, H, 64)
        Vv = torch.cat([kv_before[phys, torch.arange(pc), 1], vn[r:r + 1]], 0).double().view(pc + 1, H, 64)
        s = torch.einsum("hd,phd->hp", q[r].double().view(H, 64), K) / 8.0
        p = torch.softmax(s, -1)
        ref[r] = torch.einsum("hp,phd->hd", p, Vv).reshape(d)
    tol = 1e-5 if dtype == _lib.F32 else 2e-2
    assert (ctx.double() - ref).abs().max() < tol


@pytest.mark.parametrize("dtype", [_lib.F32, _lib.BF16, _lib.F16])
@pytest.mark.parametrize("masked", [False, True])
@pytest.mark.parametrize("S,beam", [(21, 4), (100, 4), (77, 5), (256, 12), (130, 8), (49, 1)])
def test_cross_attention_with_additive_mask(dtype, masked, S, beam):
    """S = 100 / 77: several copy stages with a partial last chunk (the K / V pipeline of the tensor-core kernel);
    beam 5: a second group of one slot per sentence."""
    B, d, H, L, layer = 3, 768, 12, 2, 1
    M = B * beam
    act = ACT[dtype]
    q = _rand((M, 3 * d), 21)
    ckv = _rand((B, S, L, 2, d), 22).to(act)
    mask = torch.ones(B, S, dtype=torch.uint8, device=DEV)
    if masked:
        mask[0, 15:] = 0
        mask[2, 3:] = 0
    ctx = torch.zeros(M, d, device=DEV, dtype=act)
    _lib.check(_lib.lib().itb_cross_attention(dtype, None, None, None, _lib.ptr(q), 3 * d, _lib.ptr(ckv), layer,
                                              _lib.ptr(mask) if masked else None, _lib.ptr(ctx), M, beam, d, H, S, L, stream()))
    torch.cuda.synchronize()
    K = ckv[:, :, layer, 0].double().view(B, S, H, 64).repeat_interleave(beam, 0)
    V = ckv[:, :, layer, 1].double().view(B, S, H, 64).repeat_interleave(beam, 0)
    s = torch.einsum("mhd,mshd->mhs", q[:, :d].double().view(M, H, 64), K) / 8.0
    s = s + ((1.0 - mask.double()) * -10000.0).repeat_interleave(beam, 0)[:, None, :]
    ref = torch.einsum("mhs,mshd->mhd", torch.softmax(s, -1), V).reshape(M, d)
    tol = 1e-5 if dtype == _lib.F32 else 2e-2
    assert (ctx.double() - ref).abs().max() < tol


# ------------------------------------------------------------------------------------------------ fused beam step
def _oracle_state(B, first_tokens):
    prefixes = torch.tensor(first_tokens).view(B, 1)
    return prefixes, torch.zeros(B, 1), None


@pytest.mark.parametrize("beam,V,n_chunk", [(4, 2000, 4), (1, 777, 1), (8, 3000, 7), (12, 5003, 3), (3, 60000, 16)])
def test_beam_step_matches_reference_semantics(beam, V, n_chunk):
    B, T, eos, pad, alpha = 5, 12, 4, 0, 0.8
    max_lens = [12, 5, 7, 12, 3]
    first = [6] * B
    h = BeamHarness(B, beam, V, T, max_lens, first, max_len=T, pad=pad, eos=eos, alpha=alpha, n_chunk=n_chunk)
    prefixes = torch.tensor(first).view(B, 1)
    scores = torch.zeros(B, 1)
    sizes = torch.zeros(B) if beam > 1 else None
    mlt = torch.tensor(max_lens)
    g = torch.Generator().manual_seed(100 + beam)
    for i in range(1, T):
        cur = prefixes.view(-1, prefixes.size(-1))
        all_fin = int(torch.any(cur == eos, 1).sum()) == cur.size(0)
        assert h.ctl()[1] == (1 if all_fin else 0)
        if all_fin:
            break
        M = B * beam
        logits = torch.randn(M, V, generator=g) * 2.0
        if i in (2, 3, 5):                       # make EOS win for some rows so rows finish at different times
            logits[torch.randperm(M, generator=g)[: M // 3], eos] += 12.0
        h.step(logits)
        rows = logits if i > 1 else logits.view(B, beam, V)[:, 0]
        logp = torch.log_softmax(rows, -1)
        prefixes, scores, sizes, _ = R.search_step(i, logp, prefixes, scores, sizes, mlt, eos, pad, beam, alpha,
                                                   topk_fn=R.lowest_index_topk)
        tok, sc, sz, anc = h.state()
        assert tok.tolist() == prefixes.tolist(), "prefixes differ at step %d" % i
        assert torch.allclose(sc, scores, rtol=1e-5, atol=1e-6), "scores differ at step %d" % i
        if beam > 1:
            assert sz.tolist() == sizes.tolist()
    assert h.ctl()[2] >= 3


@pytest.mark.parametrize("beam,V,quantum", [(4, 60000, 0.0), (4, 60000, 0.25), (4, 60000, 8.0), (8, 30000, 0.5), (12, 9000, 0.0),
                                            (1, 60000, 1.0)])
def test_beam_step_on_fused_vocabulary_partials(beam, V, quantum):
    """The 16-bit flow of the search step: phase 2 over per-chunk partials (470 chunks at V = 60000).  quantum > 0 rounds the
    logits to multiples of it, so that hundreds of chunk maxima are EQUAL: the K-th largest chunk maximum then ties with
    more chunks than the warp's queue holds and the row is ranked by the streaming form; ties resolve to the lowest index
    (reference: torch.topk on the flat (beam x V) scores, src/seq_gen.py:203)."""
    B, T, eos, pad, alpha = 3, 8, 4, 0, 0.8
    max_lens = [8, 4, 8]
    first = [6] * B
    nc = _lib.lib().itb_vocab_chunks(V)
    h = BeamHarness(B, beam, V, T, max_lens, first, max_len=T, pad=pad, eos=eos, alpha=alpha, n_chunk=nc)
    prefixes, scores = torch.tensor(first).view(B, 1), torch.zeros(B, 1)
    sizes = torch.zeros(B) if beam > 1 else None
    mlt = torch.tensor(max_lens)
    g = torch.Generator().manual_seed(300 + beam)
    for i in range(1, T):
        cur = prefixes.view(-1, prefixes.size(-1))
        if int(torch.any(cur == eos, 1).sum()) == cur.size(0):
            break
        M = B * beam
        logits = torch.randn(M, V, generator=g) * 2.0
        if quantum > 0:
            logits = torch.round(logits / quantum) * quantum
        if i == 3:
            logits[torch.randperm(M, generator=g)[: M // 3], eos] += 40.0
        h.step_fused(logits)
        rows = logits if i > 1 else logits.view(B, beam, V)[:, 0]
        logp = torch.log_softmax(rows, -1)
        prefixes, scores, sizes, _ = R.search_step(i, logp, prefixes, scores, sizes, mlt, eos, pad, beam, alpha,
                                                   topk_fn=R.lowest_index_topk)
        tok, sc, sz, anc = h.state()
        assert tok.tolist() == prefixes.tolist(), "prefixes differ at step %d" % i
        assert torch.allclose(sc, scores, rtol=1e-5, atol=2e-6), "scores differ at step %d" % i
    assert h.ctl()[2] >= 3


def test_beam_step_is_a_noop_after_termination():
    h = BeamHarness(2, 4, 1000, 6, [6, 6], [6, 6], max_len=3)
    h.step(torch.randn(8, 1000))
    h.step(torch.randn(8, 1000))
    assert h.ctl()[1] == 1 and h.ctl()[2] == 2          # max_len 3 => exactly 2 steps (seq_gen.py:131)
    before = h.state()
    h.step(torch.randn(8, 1000))
    after = h.state()
    assert h.ctl()[2] == 2 and all(torch.equal(a, b) for a, b in zip(before, after))
