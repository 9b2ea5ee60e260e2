"""Allocation-free entry point (-m gpu; collected last on purpose): itb_linear_add_ln_static against the existing pair
itb_linear (f32 epilogue) + itb_add_ln on the same inputs, called through the C ABI.  It has no counterpart in the reference
(added for callers that cannot allocate device memory); the scratch of its split-K partial outputs is a file-scope device
array of libitb.so, shared by all calls, so every case here runs it at least twice in a row with different inputs and sizes."""
import pytest
import torch

from imagetranslate_b200 import _lib

from gpu_util import ACT, DEV, stream

pytestmark = pytest.mark.gpu
EPS = 1e-12


def _inputs(dtype, M, d, K, seed):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    A = r(M, K).to(DEV).to(ACT[dtype])
    W = (r(d, K) * 0.05).to(DEV).to(ACT[dtype])
    bias, x = r(d).to(DEV), r(M, d).to(DEV)
    ln_g, ln_b = (r(d) * 0.1 + 1.0).to(DEV), (r(d) * 0.1).to(DEV)
    return A, W, bias, ln_g, ln_b, x


def _existing(dtype, impl, A, W, bias, ln_g, ln_b, x):
    """itb_linear into a caller-allocated y, then itb_add_ln over that one partial."""
    l = _lib.lib()
    M, K = A.shape
    d = W.shape[0]
    y = torch.full((M, d), float("nan"), device=DEV)
    xf, xa = x.clone(), torch.zeros(M, d, device=DEV, dtype=ACT[dtype])
    _lib.check(l.itb_linear(dtype, impl, _lib.EPI_BIAS_F32, _lib.ptr(A), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(y), M, d, K, d,
                            stream()), "linear")
    _lib.check(l.itb_add_ln(dtype, None, _lib.ptr(y), 1, _lib.ptr(ln_g), _lib.ptr(ln_b), EPS, _lib.ptr(xf), _lib.ptr(xa), M, d,
                            stream()), "add_ln")
    torch.cuda.synchronize()
    return xf, xa


def _static(dtype, impl, A, W, bias, ln_g, ln_b, x):
    l = _lib.lib()
    M, K = A.shape
    d = W.shape[0]
    xf, xa = x.clone(), torch.zeros(M, d, device=DEV, dtype=ACT[dtype])
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    _lib.check(l.itb_linear_add_ln_static(dtype, impl, _lib.ptr(A), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(ln_g), _lib.ptr(ln_b),
                                          EPS, _lib.ptr(xf), _lib.ptr(xa), M, d, K, stream()), "linear_add_ln_static")
    torch.cuda.synchronize()
    return xf, xa, free0 - torch.cuda.mem_get_info(0)[0]


CASES = [(_lib.F32, _lib.GEMM_AUTO), (_lib.F16, _lib.GEMM_TCGEN05), (_lib.BF16, _lib.GEMM_TCGEN05), (_lib.F16, _lib.GEMM_SIMT)]


@pytest.mark.parametrize("dtype,impl", CASES, ids=["f32", "f16-tcgen05", "bf16-tcgen05", "f16-simt"])
def test_static_scratch_variant_equals_linear_plus_add_ln(dtype, impl):
    """Shapes of the decoder's three LayerNorm'd projections (attention outputs K = d, feed-forward output K = 4d) at a small,
    a ragged and the largest accepted row count; back to back, so each call finds the previous call's partials in the array
    (a larger problem first, then smaller ones: stale rows and stale higher partials stay behind)."""
    limit = _lib.lib().itb_static_scratch_floats() // 4
    assert limit == 524288
    grew = []
    for n, (M, d, K) in enumerate([(244, 768, 768), (61, 768, 3072), (682, 768, 768), (7, 768, 768), (244, 768, 768)]):
        assert M * d <= limit
        args = _inputs(dtype, M, d, K, 10 + n)
        want_f, want_a = _existing(dtype, impl, *args)
        _static(dtype, impl, *args)                  # the first call of a shape may load kernel code (lazy module loading)
        got_f, got_a, used = _static(dtype, impl, *args)
        assert torch.isfinite(got_f).all()
        # the tolerances of test_embed_ln_and_add_ln: only the fp32 summation order of the split-K partials differs
        assert float((got_f - want_f).abs().max()) < 2e-5, (M, d, K)
        assert float((got_a.float() - want_a.float()).abs().max()) < (2e-5 if dtype == _lib.F32 else 3e-2), (M, d, K)
        grew.append(used)                            # ... the repeat must leave the device's free memory alone
    assert max(grew) <= 0, grew


def test_static_scratch_variant_is_deterministic_and_rejects_oversize():
    l = _lib.lib()
    args = _inputs(_lib.F16, 300, 768, 3072, 3)
    a = _static(_lib.F16, _lib.GEMM_TCGEN05, *args)
    b = _static(_lib.F16, _lib.GEMM_TCGEN05, *args)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    M, d, K = 683, 768, 768                                  # one row beyond the limit
    A, W, bias, ln_g, ln_b, x = _inputs(_lib.F16, M, d, K, 4)
    xf, xa = x.clone(), torch.zeros(M, d, device=DEV, dtype=torch.float16)
    r = l.itb_linear_add_ln_static(_lib.F16, _lib.GEMM_TCGEN05, _lib.ptr(A), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(ln_g),
                                   _lib.ptr(ln_b), EPS, _lib.ptr(xf), _lib.ptr(xa), M, d, K, stream())
    torch.cuda.synchronize()
    assert r == -2 and b"exceeds the static scratch" in l.itb_last_error()
    assert torch.equal(xf, x)                                # nothing was launched
