// Embedding + LayerNorm, residual + LayerNorm, state init, output extraction.
// HBM-bound row kernels: one CTA per row, float4 accesses, fp32 statistics.
#include <stdlib.h>
#include "common.cuh"
#include "launch.h"

// One WARP per row (d = 768 -> 24 values per lane, kept in registers), LN_ROWS rows per CTA: no shared memory, no block
// barrier; mean and variance by warp shuffles in fp32, two-pass (mean first, then centred squares) as torch.layer_norm.
#define LN_ROWS 8
#define LN_MAXV 8        // float4 per lane: d <= 32 * 4 * LN_MAXV = 1024

__device__ __forceinline__ void store4(float* p, float4 o) { *reinterpret_cast<float4*>(p) = o; }
__device__ __forceinline__ void store4(__nv_bfloat16* p, float4 o) {
    uint2 u;
    *reinterpret_cast<__nv_bfloat162*>(&u.x) = __floats2bfloat162_rn(o.x, o.y);
    *reinterpret_cast<__nv_bfloat162*>(&u.y) = __floats2bfloat162_rn(o.z, o.w);
    *reinterpret_cast<uint2*>(p) = u;
}
__device__ __forceinline__ void store4(__half* p, float4 o) {
    uint2 u;
    *reinterpret_cast<__half2*>(&u.x) = __floats2half2_rn(o.x, o.y);
    *reinterpret_cast<__half2*>(&u.y) = __floats2half2_rn(o.z, o.w);
    *reinterpret_cast<uint2*>(p) = u;
}

// NVC > 0: d = 128 * NVC is known at compile time -- the column guards fold away and the loads of a row can be issued
// together instead of one dependent round trip after the other (see add_ln_fast_kernel).
template <typename TA, int NVC = 0>
__device__ __forceinline__ void ln_row_write(float4* v, int nv, int d, const float* __restrict__ g, const float* __restrict__ b,
                                             float eps, float* __restrict__ of, TA* __restrict__ oa, int lane) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < LN_MAXV; ++i)
        if (NVC ? i < NVC : i < nv) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    const float mean = warp_sum(s) / (float)d;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < LN_MAXV; ++i)
        if (NVC ? i < NVC : i < nv) {
            float a0 = v[i].x - mean, a1 = v[i].y - mean, a2 = v[i].z - mean, a3 = v[i].w - mean;
            q += (a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3);
        }
    const float rstd = rsqrtf(warp_sum(q) / (float)d + eps);
#pragma unroll
    for (int i = 0; i < LN_MAXV; ++i)
        if (NVC ? i < NVC : i < nv) {
            const int c = (i * 32 + lane) * 4;
            if (NVC || c < d) {
                const float4 gg = *reinterpret_cast<const float4*>(g + c), bb = *reinterpret_cast<const float4*>(b + c);
                float4 o;
                o.x = (v[i].x - mean) * rstd * gg.x + bb.x;
                o.y = (v[i].y - mean) * rstd * gg.y + bb.y;
                o.z = (v[i].z - mean) * rstd * gg.z + bb.z;
                o.w = (v[i].w - mean) * rstd * gg.w + bb.w;
                *reinterpret_cast<float4*>(of + c) = o;
                store4(oa + c, o);
            }
        }
}

template <typename TA, int NVC = 0>
__global__ void __launch_bounds__(32 * LN_ROWS, NVC ? 2 : 4) embed_ln_kernel(const ItbCtl* __restrict__ ctl, const int* __restrict__ tokens,
                                                       const int* __restrict__ tgt_lang, const float* __restrict__ word,
                                                       const float* __restrict__ pos, const float* __restrict__ type,
                                                       const float* __restrict__ g, const float* __restrict__ b, float eps,
                                                       float* __restrict__ x_f32, TA* __restrict__ x_act, int M, int beam,
                                                       int d_rt, int T_max) {
    const int d = NVC ? 128 * NVC : d_rt;
    pdl_trigger();
    pdl_wait();
    const int4 c4 = *reinterpret_cast<const int4*>(ctl);          // {step, done, steps_run, cur}
    if (c4.y != 0) return;
    const int lane = threadIdx.x & 31, r = blockIdx.x * LN_ROWS + (threadIdx.x >> 5);
    if (r >= M || r >= itb_rows_active(ctl)) return;
    const int p = c4.x - 1;                          // 0-based position of the token consumed this step
    const int tok = tokens[((size_t)c4.w * M + r) * T_max + p];
    const int lang = tgt_lang[r / beam];
    const int nv = (d + 127) / 128;
    float4 v[LN_MAXV];
#pragma unroll
    for (int i = 0; i < LN_MAXV; ++i) {
        const int c = (i * 32 + lane) * 4;
        v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (NVC ? i < NVC : (i < nv && c < d)) {
            const float4 a = *reinterpret_cast<const float4*>(word + (size_t)tok * d + c);
            const float4 e = *reinterpret_cast<const float4*>(pos + (size_t)p * d + c);
            const float4 t = *reinterpret_cast<const float4*>(type + (size_t)lang * d + c);
            v[i] = make_float4(a.x + e.x + t.x, a.y + e.y + t.y, a.z + e.z + t.z, a.w + e.w + t.w);
        }
    }
    // lanes past d hold zeros, which the centred second pass must not count: handled by c < d on load and the exact
    // mean (sum / d); the variance pass subtracts the mean from padded zeros, so d must be a multiple of 128
    ln_row_write<TA, NVC>(v, nv, d, g, b, eps, x_f32 + (size_t)r * d, x_act + (size_t)r * d, lane);
}

// Encoder embeddings: every position of the padded source batch at once.  x[b*S + p] = LN(word[tok] + pos[p] + type[lang[b]])
// (HF BertEmbeddings with position_ids = arange(S); reference src/bert_seq2seq.py:103-129, token types = source language).
template <typename TA>
__global__ void __launch_bounds__(32 * LN_ROWS, 4) embed_seq_ln_kernel(const int* __restrict__ tokens, const int* __restrict__ lang,
                                                               const float* __restrict__ word, const float* __restrict__ pos,
                                                               const float* __restrict__ type, const float* __restrict__ g,
                                                               const float* __restrict__ b, float eps, float* __restrict__ x_f32,
                                                               TA* __restrict__ x_act, int rows, int S, int d) {
    const int lane = threadIdx.x & 31, r = blockIdx.x * LN_ROWS + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int tok = tokens[r], p = r % S, lg = lang[r / S];
    const int nv = (d + 127) / 128;
    float4 v[LN_MAXV];
#pragma unroll
    for (int i = 0; i < LN_MAXV; ++i) {
        const int c = (i * 32 + lane) * 4;
        v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < nv && c < d) {
            const float4 a = *reinterpret_cast<const float4*>(word + (size_t)tok * d + c);
            const float4 e = *reinterpret_cast<const float4*>(pos + (size_t)p * d + c);
            const float4 t = *reinterpret_cast<const float4*>(type + (size_t)lg * d + c);
            v[i] = make_float4(a.x + e.x + t.x, a.y + e.y + t.y, a.z + e.z + t.z, a.w + e.w + t.w);
        }
    }
    ln_row_write<TA>(v, nv, d, g, b, eps, x_f32 + (size_t)r * d, x_act + (size_t)r * d, lane);
}

template <typename TA>
__global__ void __launch_bounds__(32 * LN_ROWS, 4) add_ln_kernel(const ItbCtl* __restrict__ ctl, const float* __restrict__ y,
                                                     int n_partial, const float* __restrict__ g,
                                                     const float* __restrict__ b, float eps, float* __restrict__ x_f32,
                                                     TA* __restrict__ x_act, int M, int d) {
    pdl_trigger();
    pdl_wait();
    if (ctl != nullptr && ctl->done) return;
    const int lane = threadIdx.x & 31, r = blockIdx.x * LN_ROWS + (threadIdx.x >> 5);
    if (r >= M || r >= itb_rows_active(ctl)) return;
    const int nv = (d + 127) / 128;
    float4 v[LN_MAXV];
#pragma unroll
    for (int i = 0; i < LN_MAXV; ++i) {
        const int c = (i * 32 + lane) * 4;
        v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < nv && c < d) {
            // y = sum of the split-K partial outputs of the producing GEMM, in a fixed order (deterministic)
            float4 a = *reinterpret_cast<const float4*>(y + (size_t)r * d + c);
            for (int sp = 1; sp < n_partial; ++sp) {
                const float4 t = *reinterpret_cast<const float4*>(y + ((size_t)sp * M + r) * d + c);
                a.x += t.x; a.y += t.y; a.z += t.z; a.w += t.w;
            }
            const float4 x = *reinterpret_cast<const float4*>(x_f32 + (size_t)r * d + c);
            v[i] = make_float4(a.x + x.x, a.y + x.y, a.z + x.z, a.w + x.w);
        }
    }
    ln_row_write<TA>(v, nv, d, g, b, eps, x_f32 + (size_t)r * d, x_act + (size_t)r * d, lane);
}

// The same kernel for d = 128 * NV known at compile time, written for memory-level parallelism: the generic kernel above
// walks its float4 columns under run-time guards, so every load waits for the one before it (6 x (partials + 1) dependent
// round trips per row: ~10 us per launch at ANY row count, 30 % of a small batch's search step).  Here the row's first
// partial, its residual and the control block are requested together, the remaining partials two at a time, and the
// termination / dead-row test (which only has to precede the stores) waits for nothing else.  Same summation order.
// ld.volatile: nvcc AND ptxas keep these loads where they are written, i.e. ahead of the early-exit tests that follow them (ptxas
// would otherwise hoist the tests, and with them the control block's round trip, in front of the row's loads)
__device__ __forceinline__ float4 ld_f4_now(const float* p) {
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ int4 ld_i4_now(const void* p) {
    int4 v;
    asm volatile("ld.volatile.global.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

template <typename TA, int NV>
__global__ void __launch_bounds__(32 * LN_ROWS, 2) add_ln_fast_kernel(const ItbCtl* __restrict__ ctl, const float* __restrict__ y,
                                                          int n_partial, const float* __restrict__ g,
                                                          const float* __restrict__ b, float eps, float* __restrict__ x_f32,
                                                          TA* __restrict__ x_act, int M) {
    pdl_trigger();
    pdl_wait();
    constexpr int d = 128 * NV;
    const int lane = threadIdx.x & 31, r = blockIdx.x * LN_ROWS + (threadIdx.x >> 5);
    if (r >= M) return;
    int4 c0 = make_int4(0, 0, 0, 0), c1 = make_int4(0, 0, 0, 0);
    if (ctl != nullptr) {
        c0 = ld_i4_now(ctl);                                       // {step, done, steps_run, cur}
        c1 = ld_i4_now(reinterpret_cast<const int4*>(ctl) + 1);    // {fin_rows, sent_done, max_len, rows_active}
    }
    const float* yr = y + (size_t)r * d + lane * 4;
    float* xr = x_f32 + (size_t)r * d + lane * 4;
    float4 a[NV], xv[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) a[i] = ld_f4_now(yr + i * 128);
#pragma unroll
    for (int i = 0; i < NV; ++i) xv[i] = ld_f4_now(xr + i * 128);
    if (c0.y != 0 || (c1.w > 0 && r >= c1.w)) return;            // search over / row of a finished group (itb_rows_active)
    const size_t ps = (size_t)M * d;
    int sp = 1;
    for (; sp + 1 < n_partial; sp += 2) {
        float4 t0[NV], t1[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) t0[i] = *reinterpret_cast<const float4*>(yr + sp * ps + i * 128);
#pragma unroll
        for (int i = 0; i < NV; ++i) t1[i] = *reinterpret_cast<const float4*>(yr + (sp + 1) * ps + i * 128);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            a[i].x += t0[i].x; a[i].y += t0[i].y; a[i].z += t0[i].z; a[i].w += t0[i].w;
            a[i].x += t1[i].x; a[i].y += t1[i].y; a[i].z += t1[i].z; a[i].w += t1[i].w;
        }
    }
    if (sp < n_partial) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const float4 t = *reinterpret_cast<const float4*>(yr + sp * ps + i * 128);
            a[i].x += t.x; a[i].y += t.y; a[i].z += t.z; a[i].w += t.w;
        }
    }
    // LayerNorm as ln_row_write (two-pass statistics, same expression order), all NV columns live
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        a[i] = make_float4(a[i].x + xv[i].x, a[i].y + xv[i].y, a[i].z + xv[i].z, a[i].w + xv[i].w);
        s += (a[i].x + a[i].y) + (a[i].z + a[i].w);
    }
    const float mean = warp_sum(s) / (float)d;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const float a0 = a[i].x - mean, a1 = a[i].y - mean, a2 = a[i].z - mean, a3 = a[i].w - mean;
        q += (a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3);
    }
    const float rstd = rsqrtf(warp_sum(q) / (float)d + eps);
    TA* ar = x_act + (size_t)r * d + lane * 4;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const float4 gg = *reinterpret_cast<const float4*>(g + lane * 4 + i * 128), bb = *reinterpret_cast<const float4*>(b + lane * 4 + i * 128);
        float4 o;
        o.x = (a[i].x - mean) * rstd * gg.x + bb.x;
        o.y = (a[i].y - mean) * rstd * gg.y + bb.y;
        o.z = (a[i].z - mean) * rstd * gg.z + bb.z;
        o.w = (a[i].w - mean) * rstd * gg.w + bb.w;
        *reinterpret_cast<float4*>(xr + i * 128) = o;
        store4(ar + i * 128, o);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Two-stream decoding (reference src/seq_gen.py:167-188).  The second decoder pass (object-feature stream through
// `obj_decoder`, or the image memory through the same decoder in the text+image mode) runs on the same buffers as the
// first; the first stream's last state is parked in x_first meanwhile, and the two are mixed by a learned sigmoid gate:
//   h = sigmoid(gate + 1e-7) * h_first + (1 - sigmoid(gate + 1e-7)) * h_second
__global__ void save_state_kernel(const ItbCtl* __restrict__ ctl, const float4* __restrict__ x, float4* __restrict__ dst, int n4) {
    pdl_trigger();
    pdl_wait();
    if (ctl != nullptr && ctl->done) return;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x) dst[i] = x[i];
}

template <typename TA>
__global__ void gate_mix_kernel(const ItbCtl* __restrict__ ctl, const float* __restrict__ x_first, const float* __restrict__ gate,
                                float eps_gate, float* __restrict__ x_f32, TA* __restrict__ x_act, int M, int d) {
    pdl_trigger();
    pdl_wait();
    if (ctl != nullptr && ctl->done) return;
    const int n4 = M * d / 4, d4 = d / 4;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x) {
        const float4 a = reinterpret_cast<const float4*>(x_first)[i], b = reinterpret_cast<const float4*>(x_f32)[i];
        const float4 g = reinterpret_cast<const float4*>(gate)[i % d4];
        const float s0 = 1.f / (1.f + expf(-(g.x + eps_gate))), s1 = 1.f / (1.f + expf(-(g.y + eps_gate)));
        const float s2 = 1.f / (1.f + expf(-(g.z + eps_gate))), s3 = 1.f / (1.f + expf(-(g.w + eps_gate)));
        float4 o;
        o.x = s0 * a.x + (1.f - s0) * b.x;
        o.y = s1 * a.y + (1.f - s1) * b.y;
        o.z = s2 * a.z + (1.f - s2) * b.z;
        o.w = s3 * a.w + (1.f - s3) * b.w;
        reinterpret_cast<float4*>(x_f32)[i] = o;
        store4(x_act + (size_t)i * 4, o);
    }
}

// Lexical proposals (reference src/seq2seq.py:110-144 `attend_proposal`, called on the last state at src/seq_gen.py:191-192).
// Per decoder row (sentence b = row / beam): scores_p = h . E[prop[b][p]] (no scaling), probs = softmax_p(scores),
// ctx = sum_p probs_p E[prop[b][p]]; ctx := 1e-8 when every proposal of the sentence is PAD;
// h := LayerNorm(sigmoid(lexical_gate + 1e-8) * h + (1 - sigmoid(..)) * ctx).
// The reference's pad mask `attend_scores[attend_mask].fill_(-10000.0)` (:131) fills a COPY (boolean-mask indexing), so
// PAD proposals are attended with the PAD row of the embedding table like any other: reproduced, not "fixed".
#define AP_THREADS 256
#define AP_MAXP 256
template <typename TA>
__global__ void __launch_bounds__(AP_THREADS) attend_proposal_kernel(
    const ItbCtl* __restrict__ ctl, const int* __restrict__ proposals, int n_prop, int pad, const float* __restrict__ emb,
    const float* __restrict__ lex_gate, const float* __restrict__ g, const float* __restrict__ b, float eps,
    float* __restrict__ x_f32, TA* __restrict__ x_act, int beam, int d) {
    pdl_trigger();
    pdl_wait();
    if (ctl != nullptr && ctl->done) return;
    extern __shared__ float ap_smem[];
    float* xs = ap_smem;                 // [d]
    float* sc = ap_smem + d;             // [AP_MAXP]
    __shared__ float red[32];
    __shared__ int all_pad;
    const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int* prop = proposals + (size_t)(r / beam) * n_prop;
    float* xr = x_f32 + (size_t)r * d;
    if (tid == 0) all_pad = 1;
    for (int c = tid; c < d; c += AP_THREADS) xs[c] = xr[c];
    __syncthreads();
    for (int p = w; p < n_prop; p += AP_THREADS / 32) {
        const int id = prop[p];
        const float* e = emb + (size_t)id * d;
        float a = 0.f;
        for (int c = lane; c < d; c += 32) a = fmaf(xs[c], e[c], a);
        a = warp_sum(a);
        if (lane == 0) {
            sc[p] = a;
            if (id != pad) all_pad = 0;
        }
    }
    __syncthreads();
    float mx = -3.0e38f;
    for (int p = 0; p < n_prop; ++p) mx = fmaxf(mx, sc[p]);
    float den = 0.f;
    for (int p = 0; p < n_prop; ++p) den += expf(sc[p] - mx);
    const float inv = 1.f / den;
    const bool ap = all_pad != 0;
    // mixed value per column, then LayerNorm (two-pass, fp32)
    float loc[4];                        // d <= 4 * AP_THREADS
    float s = 0.f;
    int k = 0;
    for (int c = tid; c < d; c += AP_THREADS, ++k) {
        float ctx = 0.f;
        if (ap) ctx = 1e-8f;
        else
            for (int p = 0; p < n_prop; ++p) ctx = fmaf(expf(sc[p] - mx) * inv, emb[(size_t)prop[p] * d + c], ctx);
        const float sg = 1.f / (1.f + expf(-(lex_gate[c] + 1e-8f)));
        loc[k] = sg * xs[c] + (1.f - sg) * ctx;
        s += loc[k];
    }
    const float mean = block_sum(s, red) / (float)d;
    float q = 0.f;
    k = 0;
    for (int c = tid; c < d; c += AP_THREADS, ++k) q += (loc[k] - mean) * (loc[k] - mean);
    const float rstd = rsqrtf(block_sum(q, red) / (float)d + eps);
    k = 0;
    for (int c = tid; c < d; c += AP_THREADS, ++k) {
        const float o = (loc[k] - mean) * rstd * g[c] + b[c];
        xr[c] = o;
        x_act[(size_t)r * d + c] = from_f32<TA>(o);
    }
}

// Physical reorder of the self-attention KV cache by back-pointer: dst[l][r][p] = src[l][src_row[r]][p] for p < n_pos.
// This is what a direct translation of the reference's `top_beam_outputs.gather(1, beam_indices)` (src/seq_gen.py:217-219)
// implies for a KV cache.  The decode path does NOT use it -- the ancestor table replaces it -- it exists so that the two
// variants can be measured against each other (tools/kv_reorder_bench.py, profiles/).
__global__ void kv_gather_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, const int* __restrict__ src_row,
                                 int M, size_t row_units, int n_units) {
    const int r = blockIdx.x, l = blockIdx.y;
    const uint4* s = src + ((size_t)l * M + src_row[r]) * row_units;
    uint4* d = dst + ((size_t)l * M + r) * row_units;
    for (int i = threadIdx.x; i < n_units; i += blockDim.x) d[i] = s[i];
}

// seq_gen.py:109-123: prefix = first token in every slot, scores 0, sizes 0; ancestors = own slot.
__global__ void init_state_kernel(ItbCtl* ctl, ItbGrp* grp, int* sent_done, const int* __restrict__ sent_grp,
                                  const int* __restrict__ grp_max_len, int G, int* tokens, unsigned char* anc,
                                  float* scores, float* sizes, unsigned char* fin, int* sent_counter,
                                  const int* __restrict__ first_tokens, int B, int beam, int T_max, int eos) {
    const int M = B * beam;
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = gridDim.x * blockDim.x;
    for (size_t i = tid; i < (size_t)2 * M * T_max; i += n) {
        int r = (int)((i / T_max) % M);
        int p = (int)(i % T_max);
        tokens[i] = (p == 0) ? first_tokens[r / beam] : 0;
        anc[i] = (unsigned char)(r % beam);
    }
    for (int i = tid; i < 2 * M; i += n) {
        scores[i] = 0.f;
        sizes[i] = 0.f;
        fin[i] = (first_tokens[(i % M) / beam] == eos) ? 1 : 0;
    }
    for (int i = tid; i < B; i += n) sent_counter[i] = 0;
    if (blockIdx.x == 0) {
        // group bookkeeping by one CTA (G and B are small)
        __shared__ int all_done;
        if (threadIdx.x == 0) all_done = 1;
        __syncthreads();
        for (int g = threadIdx.x; g < G; g += blockDim.x) {
            int rows = 0, fin_rows = 0;
            for (int b = 0; b < B; ++b)
                if (sent_grp[b] == g) { rows += beam; fin_rows += (first_tokens[b] == eos) ? beam : 0; }
            const int ml = grp_max_len[g];
            // seq_gen.py:131-136: loop body runs for i in [1, max_len) unless every row already holds EOS
            const int done = (ml <= 1 || fin_rows == rows) ? 1 : 0;
            grp[g].done = done; grp[g].steps_run = 0; grp[g].max_len = ml; grp[g].fin_rows = 0;
            grp[g].n_rows = rows; grp[g].final_cur = 0; grp[g].pad0_ = 0; grp[g].pad1_ = 0;
            if (!done) atomicAnd(&all_done, 0);
        }
        __syncthreads();
        for (int b = threadIdx.x; b < B; b += blockDim.x) sent_done[b] = grp[sent_grp[b]].done;
        if (threadIdx.x == 0) {
            int ml = 0;
            for (int g = 0; g < G; ++g) ml = max(ml, grp_max_len[g]);
            ctl->step = 1;
            ctl->steps_run = 0;
            ctl->cur = 0;
            ctl->fin_rows = 0;
            ctl->sent_done = 0;
            ctl->max_len = ml;
            ctl->rows_active = M;
            ctl->done = all_done;
        }
    }
}

// seq_gen.py:229-231 + :6-24 : slot 0, cut before first EOS else at max_lens (clamped to the row length).
__global__ void finalize_kernel(const ItbCtl* __restrict__ ctl, const ItbGrp* __restrict__ grp,
                                const int* __restrict__ sent_grp, const int* __restrict__ tokens,
                                const int* __restrict__ max_lens, int eos, int* __restrict__ out_tokens,
                                int* __restrict__ out_len, int* __restrict__ out_len_full, int B, int beam, int T_max) {
    const int b = blockIdx.x;
    const int M = B * beam;
    const ItbGrp g = grp[sent_grp[b]];
    const int n_tok = g.steps_run + 1;
    const int cur = g.done ? g.final_cur : ctl->cur;
    const int* src = tokens + ((size_t)cur * M + (size_t)b * beam) * T_max;
    __shared__ int first_eos;
    if (threadIdx.x == 0) first_eos = 0x7fffffff;
    __syncthreads();
    for (int p = threadIdx.x; p < T_max; p += blockDim.x) {
        int t = (p < n_tok) ? src[p] : 0;
        out_tokens[(size_t)b * T_max + p] = t;
        if (p < n_tok && t == eos) atomicMin(&first_eos, p);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int len = (first_eos != 0x7fffffff) ? first_eos : min(max_lens[b], n_tok);
        out_len[b] = len;
        out_len_full[b] = n_tok;
    }
}

extern "C" {

int itb_embed_ln(int dtype, const itb_ctl* ctl, const int* tokens, const int* tgt_lang, const float* word,
                 const float* pos, const float* type, const float* g, const float* b, float eps, float* x_f32,
                 void* x_act, int M, int beam, int d, int T_max, void* stream) {
    ITB_REQUIRE(d % 128 == 0 && d <= 1024 && M > 0, "embed_ln: d must be a multiple of 128, at most 1024");
    const size_t smem = 0;
    const dim3 lgrid((M + LN_ROWS - 1) / LN_ROWS), lblock(32 * LN_ROWS);
    cudaStream_t s = (cudaStream_t)stream;
    const ItbCtl* c = reinterpret_cast<const ItbCtl*>(ctl);
    cudaError_t e;
    if (d == 768 && dtype == ITB_BF16)
        e = itb_launch(embed_ln_kernel<__nv_bfloat16, 6>, lgrid, lblock, smem, s, c, tokens, tgt_lang, word, pos, type, g, b,
                       eps, x_f32, (__nv_bfloat16*)x_act, M, beam, d, T_max);
    else if (d == 768 && dtype == ITB_F16)
        e = itb_launch(embed_ln_kernel<__half, 6>, lgrid, lblock, smem, s, c, tokens, tgt_lang, word, pos, type, g, b, eps,
                       x_f32, (__half*)x_act, M, beam, d, T_max);
    else if (dtype == ITB_BF16)
        e = itb_launch(embed_ln_kernel<__nv_bfloat16>, lgrid, lblock, smem, s, c, tokens, tgt_lang, word, pos, type, g, b,
                       eps, x_f32, (__nv_bfloat16*)x_act, M, beam, d, T_max);
    else if (dtype == ITB_F16)
        e = itb_launch(embed_ln_kernel<__half>, lgrid, lblock, smem, s, c, tokens, tgt_lang, word, pos, type, g, b, eps,
                       x_f32, (__half*)x_act, M, beam, d, T_max);
    else
        e = itb_launch(embed_ln_kernel<float>, lgrid, lblock, smem, s, c, tokens, tgt_lang, word, pos, type, g, b, eps,
                       x_f32, (float*)x_act, M, beam, d, T_max);
    if (e != cudaSuccess) { itb_set_error("embed_ln: launch failed: %s", cudaGetErrorString(e)); return -1; }
    return itb_check_launch("embed_ln");
}

int itb_embed_seq_ln(int dtype, const int* tokens, const int* lang, const float* word, const float* pos, const float* type,
                     const float* g, const float* b, float eps, float* x_f32, void* x_act, int B, int S, int d, void* stream) {
    ITB_REQUIRE(d % 128 == 0 && d <= 1024 && B > 0 && S > 0, "embed_seq_ln: d must be a multiple of 128, at most 1024");
    const int rows = B * S;
    const dim3 lgrid((rows + LN_ROWS - 1) / LN_ROWS), lblock(32 * LN_ROWS);
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == ITB_BF16)
        embed_seq_ln_kernel<__nv_bfloat16><<<lgrid, lblock, 0, s>>>(tokens, lang, word, pos, type, g, b, eps, x_f32, (__nv_bfloat16*)x_act, rows, S, d);
    else if (dtype == ITB_F16)
        embed_seq_ln_kernel<__half><<<lgrid, lblock, 0, s>>>(tokens, lang, word, pos, type, g, b, eps, x_f32, (__half*)x_act, rows, S, d);
    else
        embed_seq_ln_kernel<float><<<lgrid, lblock, 0, s>>>(tokens, lang, word, pos, type, g, b, eps, x_f32, (float*)x_act, rows, S, d);
    return itb_check_launch("embed_seq_ln");
}

int itb_add_ln(int dtype, const itb_ctl* ctl, const float* y, int n_partial, const float* g, const float* b, float eps,
               float* x_f32, void* x_act, int M, int d, void* stream) {
    ITB_REQUIRE(d % 128 == 0 && d <= 1024 && M > 0, "add_ln: d must be a multiple of 128, at most 1024");
    ITB_REQUIRE(n_partial >= 1, "add_ln: n_partial must be >= 1");
    const size_t smem = 0;
    const dim3 lgrid((M + LN_ROWS - 1) / LN_ROWS), lblock(32 * LN_ROWS);
    cudaStream_t s = (cudaStream_t)stream;
    const ItbCtl* c = reinterpret_cast<const ItbCtl*>(ctl);
    cudaError_t e;
    static int fast = -1;                        // ITB_LN_FAST=0: the generic kernel at every width (A/B timing)
    if (fast < 0) {
        const char* ev = getenv("ITB_LN_FAST");
        fast = (ev != nullptr && ev[0] == '0') ? 0 : 1;
    }
    if (fast != 0 && d == 768) {
        if (dtype == ITB_BF16)
            e = itb_launch(add_ln_fast_kernel<__nv_bfloat16, 6>, lgrid, lblock, smem, s, c, y, n_partial, g, b, eps, x_f32,
                           (__nv_bfloat16*)x_act, M);
        else if (dtype == ITB_F16)
            e = itb_launch(add_ln_fast_kernel<__half, 6>, lgrid, lblock, smem, s, c, y, n_partial, g, b, eps, x_f32, (__half*)x_act, M);
        else
            e = itb_launch(add_ln_fast_kernel<float, 6>, lgrid, lblock, smem, s, c, y, n_partial, g, b, eps, x_f32, (float*)x_act, M);
    } else if (dtype == ITB_BF16)
        e = itb_launch(add_ln_kernel<__nv_bfloat16>, lgrid, lblock, smem, s, c, y, n_partial, g, b, eps, x_f32,
                       (__nv_bfloat16*)x_act, M, d);
    else if (dtype == ITB_F16)
        e = itb_launch(add_ln_kernel<__half>, lgrid, lblock, smem, s, c, y, n_partial, g, b, eps, x_f32, (__half*)x_act, M, d);
    else
        e = itb_launch(add_ln_kernel<float>, lgrid, lblock, smem, s, c, y, n_partial, g, b, eps, x_f32, (float*)x_act, M, d);
    if (e != cudaSuccess) { itb_set_error("add_ln: launch failed: %s", cudaGetErrorString(e)); return -1; }
    return itb_check_launch("add_ln");
}

int itb_save_state(const itb_ctl* ctl, const float* x_f32, float* x_first, int M, int d, void* stream) {
    ITB_REQUIRE(d % 4 == 0 && M > 0, "save_state: d must be a multiple of 4");
    const int n4 = M * d / 4;
    cudaError_t e = itb_launch(save_state_kernel, dim3((n4 + 255) / 256 < 1184 ? (n4 + 255) / 256 : 1184), dim3(256), (size_t)0,
                               (cudaStream_t)stream, reinterpret_cast<const ItbCtl*>(ctl), reinterpret_cast<const float4*>(x_f32),
                               reinterpret_cast<float4*>(x_first), n4);
    if (e != cudaSuccess) { itb_set_error("save_state: launch failed: %s", cudaGetErrorString(e)); return -1; }
    return itb_check_launch("save_state");
}

int itb_gate_mix(int dtype, const itb_ctl* ctl, const float* x_first, const float* gate, float gate_eps, float* x_f32, void* x_act,
                 int M, int d, void* stream) {
    ITB_REQUIRE(d % 4 == 0 && M > 0, "gate_mix: d must be a multiple of 4");
    const int n4 = M * d / 4;
    const dim3 grid((n4 + 255) / 256 < 1184 ? (n4 + 255) / 256 : 1184), block(256);
    cudaStream_t s = (cudaStream_t)stream;
    const ItbCtl* c = reinterpret_cast<const ItbCtl*>(ctl);
    cudaError_t e;
    if (dtype == ITB_BF16) e = itb_launch(gate_mix_kernel<__nv_bfloat16>, grid, block, (size_t)0, s, c, x_first, gate, gate_eps, x_f32, (__nv_bfloat16*)x_act, M, d);
    else if (dtype == ITB_F16) e = itb_launch(gate_mix_kernel<__half>, grid, block, (size_t)0, s, c, x_first, gate, gate_eps, x_f32, (__half*)x_act, M, d);
    else e = itb_launch(gate_mix_kernel<float>, grid, block, (size_t)0, s, c, x_first, gate, gate_eps, x_f32, (float*)x_act, M, d);
    if (e != cudaSuccess) { itb_set_error("gate_mix: launch failed: %s", cudaGetErrorString(e)); return -1; }
    return itb_check_launch("gate_mix");
}

int itb_attend_proposal(int dtype, const itb_ctl* ctl, const int* proposals, int n_prop, int pad_id, const float* prop_emb,
                        const float* lex_gate, const float* ln_g, const float* ln_b, float eps, float* x_f32, void* x_act,
                        int M, int beam, int d, void* stream) {
    ITB_REQUIRE(n_prop >= 1 && n_prop <= AP_MAXP, "attend_proposal: 1..256 proposals per sentence");
    ITB_REQUIRE(d <= 4 * AP_THREADS && M > 0 && beam >= 1, "attend_proposal: hidden size above 1024");
    const dim3 grid(M), block(AP_THREADS);
    const size_t smem = (size_t)(d + AP_MAXP) * sizeof(float);
    cudaStream_t s = (cudaStream_t)stream;
    const ItbCtl* c = reinterpret_cast<const ItbCtl*>(ctl);
    cudaError_t e;
    if (dtype == ITB_BF16) e = itb_launch(attend_proposal_kernel<__nv_bfloat16>, grid, block, smem, s, c, proposals, n_prop, pad_id, prop_emb, lex_gate, ln_g, ln_b, eps, x_f32, (__nv_bfloat16*)x_act, beam, d);
    else if (dtype == ITB_F16) e = itb_launch(attend_proposal_kernel<__half>, grid, block, smem, s, c, proposals, n_prop, pad_id, prop_emb, lex_gate, ln_g, ln_b, eps, x_f32, (__half*)x_act, beam, d);
    else e = itb_launch(attend_proposal_kernel<float>, grid, block, smem, s, c, proposals, n_prop, pad_id, prop_emb, lex_gate, ln_g, ln_b, eps, x_f32, (float*)x_act, beam, d);
    if (e != cudaSuccess) { itb_set_error("attend_proposal: launch failed: %s", cudaGetErrorString(e)); return -1; }
    return itb_check_launch("attend_proposal");
}

int itb_kv_gather(int dtype, const void* kv_src, void* kv_dst, const int* src_row, int M, int T_max, int n_pos, int d, int L,
                  void* stream) {
    ITB_REQUIRE(M > 0 && L > 0 && n_pos >= 0 && n_pos <= T_max && d % 8 == 0, "kv_gather: bad sizes");
    const size_t esz = dtype == ITB_F32 ? 4 : 2;
    const size_t row_units = (size_t)T_max * 2 * d * esz / 16;
    const int n_units = (int)((size_t)n_pos * 2 * d * esz / 16);
    if (n_units == 0) return 0;
    kv_gather_kernel<<<dim3(M, L), 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const uint4*>(kv_src), reinterpret_cast<uint4*>(kv_dst),
                                                                    src_row, M, row_units, n_units);
    return itb_check_launch("kv_gather");
}

int itb_init_state(const itb_config* c, const int* first_tokens, const int* grp_max_len, void* stream) {
    ITB_REQUIRE(c->G >= 1 && c->grp != nullptr && c->sent_grp != nullptr && c->sent_done != nullptr, "init_state: group arrays missing");
    init_state_kernel<<<64, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<ItbCtl*>(c->ctl), reinterpret_cast<ItbGrp*>(c->grp), c->sent_done, c->sent_grp, grp_max_len, c->G,
        c->tokens, c->anc, c->scores, c->sizes, c->fin, c->sent_counter, first_tokens, c->B, c->beam, c->T_max, c->eos_id);
    return itb_check_launch("init_state");
}

int itb_finalize(const itb_ctl* ctl, const itb_grp* grp, const int* sent_grp, const int* tokens, const int* max_lens,
                 int eos, int* out_tokens, int* out_len, int* out_len_full, int B, int beam, int T_max, void* stream) {
    finalize_kernel<<<B, 128, 0, (cudaStream_t)stream>>>(reinterpret_cast<const ItbCtl*>(ctl),
                                                         reinterpret_cast<const ItbGrp*>(grp), sent_grp, tokens, max_lens,
                                                         eos, out_tokens, out_len, out_len_full, B, beam, T_max);
    return itb_check_launch("finalize");
}

}  // extern "C"
