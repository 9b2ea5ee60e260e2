// Single-query attention for the incremental decoder step (head_dim fixed at 64 = 768/12,
// reference src/lm_config.py:8,13; HF BertSelfAttention).
//
// One CTA (4 warps) per (sentence, head, group of <=4 beam slots).  The beam slots of a sentence share K/V rows:
//   * cross-attention: all of the memory K/V, so each K/V byte is fetched once per group of slots;
//   * self-attention: the slots' hypotheses descend from a common ancestor, so through the ancestor table written
//     by the beam kernel they address the SAME physical cache rows for every position before their lineages split
//     (if two slots share the row of position p they share all earlier rows).  Positions [0,c) are therefore read
//     once for the whole group and only the divergent tail [c,n) per slot -- the K/V traffic of a beam-4 step is
//     close to that of ONE hypothesis, and a beam reorder never copies K/V.
// Positions are dealt to the warps in runs of 16; inside a warp 8 lanes x 16 B cover one 64-wide head row, 4 positions
// side by side, 4 such loads in flight per lane:
//   phase A  scores  q.k/8 (+ additive -10000 source mask) -> shared memory
//   phase B  softmax statistics in fp32: exp(x - max) / sum, as torch.softmax
//   phase C  sum_p e_p * V[p], reduced over lanes and warps, normalised, stored in the activation type.
// Self-attention appends the rows' new K,V to the cache first.  HBM-bound: 2*d*sizeof(T) bytes per (hypothesis
// lineage, position, layer) for self-attention, per (sentence, position, layer) for cross-attention.
#include <stdlib.h>
#include "common.cuh"
#include "launch.h"

#define DH 64
#define NQ_MAX 4
#define AT_WARPS 4
#define AT_THREADS (AT_WARPS * 32)

template <typename T> __device__ __forceinline__ void store8(T* p, const float* v);
template <> __device__ __forceinline__ void store8<float>(float* p, const float* v) {
    reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
}
template <> __device__ __forceinline__ void store8<__nv_bfloat16>(__nv_bfloat16* p, const float* v) {
    uint4 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int e = 0; e < 4; ++e) h[e] = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
    *reinterpret_cast<uint4*>(p) = u;
}
template <> __device__ __forceinline__ void store8<__half>(__half* p, const float* v) {
    uint4 u;
    __half2* h = reinterpret_cast<__half2*>(&u);
#pragma unroll
    for (int e = 0; e < 4; ++e) h[e] = __floats2half2_rn(v[2 * e], v[2 * e + 1]);
    *reinterpret_cast<uint4*>(p) = u;
}

template <typename T> struct Raw8;                       // 8 consecutive head dims of one K/V row
template <> struct Raw8<float> {
    float4 a, b;
    __device__ __forceinline__ void load(const float* p) {
        a = reinterpret_cast<const float4*>(p)[0];
        b = reinterpret_cast<const float4*>(p)[1];
    }
    __device__ __forceinline__ void unpack(float* o) const {
        o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
    }
};
template <> struct Raw8<__nv_bfloat16> {
    uint4 u;
    __device__ __forceinline__ void load(const __nv_bfloat16* p) { u = *reinterpret_cast<const uint4*>(p); }
    __device__ __forceinline__ void unpack(float* o) const {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
        for (int e = 0; e < 4; ++e) { float2 f = __bfloat1622float2(h[e]); o[2 * e] = f.x; o[2 * e + 1] = f.y; }
    }
};
template <> struct Raw8<__half> {
    uint4 u;
    __device__ __forceinline__ void load(const __half* p) { u = *reinterpret_cast<const uint4*>(p); }
    __device__ __forceinline__ void unpack(float* o) const {
        const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
        for (int e = 0; e < 4; ++e) { float2 f = __half22float2(h[e]); o[2 * e] = f.x; o[2 * e + 1] = f.y; }
    }
};

struct AttnArgs {
    const ItbCtl* ctl;
    const int* sent_done;                // [B] 1 = the sentence's group has terminated (rows are skipped); may be null
    const int* mem_len;                  // cross: [B] memory positions of the sentence (<= S); null = S
    const float* q; int ldq;             // f32 queries, row stride ldq (self: the qkv buffer, k at +d, v at +2d)
    void* kv;                            // self: cache of this layer [M][T_max][2][d]; cross: [B][S][L][2][d]
    const unsigned char* anc;            // self: [2][M][T_max]
    const unsigned char* mask;           // cross: [B][S] or null
    void* ctx;                           // act [M][d]
    int M, beam, d, T_max, S, L, layer;
};



// NQ = beam slots handled per CTA (4, or 1 for greedy search).  grid = (sentences, heads, ceil(beam / NQ)).
template <typename T, int NQ, bool SELF>
__global__ void __launch_bounds__(AT_THREADS, 6) attn_kernel(AttnArgs a) {
    extern __shared__ __align__(16) float sm[];
    __shared__ int s_c;
    pdl_trigger();
    pdl_wait();
    // one round trip for all control words: {step, done, steps_run, cur} and the sentence's group flag
    const int b = blockIdx.x;
    int4 c4 = make_int4(1, 0, 0, 0);
    if (a.ctl != nullptr) c4 = *reinterpret_cast<const int4*>(a.ctl);
    const int sdone = (a.sent_done != nullptr) ? a.sent_done[b] : 0;
    if (c4.y != 0 || sdone != 0) return;
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, g = lane >> 3, j = lane & 7;
    const int h = blockIdx.y, d = a.d, q0 = blockIdx.z * NQ;
    const int row0 = b * a.beam + q0, nq = min(NQ, a.beam - q0);
    const int n = SELF ? c4.x : ((a.mem_len != nullptr) ? a.mem_len[b] : a.S);    // self: positions 0..step-1
    const int npad = (n + 3) & ~3;
    float* sc = sm;                                      // [NQ][npad] scores, then exp(score - max)
    float* pv = sc + NQ * npad;                          // [AT_WARPS][NQ][DH]
    float* qs = pv + AT_WARPS * NQ * DH;                 // [NQ][DH] pre-scaled queries
    float* inv = qs + NQ * DH;                           // [NQ]
    unsigned char* s_anc = reinterpret_cast<unsigned char*>(inv + NQ);     // SELF: [NQ][n] physical beam slot per position

    // second round trip: queries, the rows' new K,V, the ancestor rows -- all loads issued before any store
    constexpr int QPT = (NQ * DH + AT_THREADS - 1) / AT_THREADS;           // query elements per thread
    constexpr int APT = (NQ * 2 * DH + AT_THREADS - 1) / AT_THREADS;       // appended K,V elements per thread
    float qv[QPT], av[APT];
#pragma unroll
    for (int x = 0; x < QPT; ++x) {
        const int i = tid + x * AT_THREADS;
        const int qi = i / DH, r = row0 + min(qi, nq - 1);
        qv[x] = (i < NQ * DH) ? a.q[(size_t)r * a.ldq + h * DH + (i % DH)] : 0.f;
    }
    int c = n;                                           // positions [0,c) are shared by the nq slots
    const T* kbase;
    size_t pos_stride;
    if (SELF) {
        const int pc = n - 1;
#pragma unroll
        for (int x = 0; x < APT; ++x) {
            const int i = tid + x * AT_THREADS;
            const int qi = i / (2 * DH), kvsel = (i / DH) & 1, e = i % DH, r = row0 + qi;
            av[x] = (i < nq * 2 * DH) ? a.q[(size_t)r * a.ldq + (1 + kvsel) * d + h * DH + e] : 0.f;
        }
        for (int i = tid; i < NQ * pc; i += AT_THREADS) {
            const int qi = i / pc, p = i % pc;
            const unsigned char* arow = a.anc + ((size_t)c4.w * a.M + row0 + min(qi, nq - 1)) * a.T_max;
            s_anc[qi * n + p] = arow[p];
        }
        // append the rows' K,V (rounded to the cache type) at position pc, then treat it like any cached position
#pragma unroll
        for (int x = 0; x < APT; ++x) {
            const int i = tid + x * AT_THREADS;
            const int qi = i / (2 * DH), kvsel = (i / DH) & 1, e = i % DH, r = row0 + qi;
            if (i < nq * 2 * DH)
                reinterpret_cast<T*>(a.kv)[(((size_t)r * a.T_max + pc) * 2 + kvsel) * d + h * DH + e] = from_f32<T>(av[x]);
        }
        if (tid < NQ) s_anc[tid * n + pc] = (unsigned char)((q0 + min(tid, nq - 1)) % a.beam);
        if (tid == 0) s_c = pc;
    }
#pragma unroll
    for (int x = 0; x < QPT; ++x) {
        const int i = tid + x * AT_THREADS;
        if (i < NQ * DH) qs[i] = qv[x] * 0.125f;         // pre-scaled by 1/sqrt(64) (exact: power of two)
    }
    __syncthreads();
    if (SELF) {
        const int pc = n - 1;
        if (NQ > 1) {
            int cmin = pc;
            for (int p = tid; p < pc; p += AT_THREADS) {
                bool same = true;
#pragma unroll
                for (int qi = 1; qi < NQ; ++qi) same = same && (s_anc[qi * n + p] == s_anc[p]);
                if (!same) { cmin = p; break; }          // p ascending per thread: first mismatch is this thread's minimum
            }
            if (cmin < pc) atomicMin(&s_c, cmin);
            __syncthreads();
            c = s_c;
        }
        kbase = reinterpret_cast<const T*>(a.kv) + h * DH + j * 8;
        pos_stride = (size_t)2 * d;
    } else {
        kbase = reinterpret_cast<const T*>(a.kv) + ((size_t)b * a.S * a.L + a.layer) * 2 * d + h * DH + j * 8;
        pos_stride = (size_t)a.L * 2 * d;
    }
    const int slot_base = b * a.beam;
    // K row (this lane's 8 dims) of position p as seen by slot qi; V row = K row + d
    auto kaddr = [&](int qi, int p) -> const T* {
        if (SELF) return kbase + ((size_t)(slot_base + s_anc[qi * n + p]) * a.T_max + p) * pos_stride;
        return kbase + (size_t)p * pos_stride;
    };
    float qr[NQ][8];
#pragma unroll
    for (int qi = 0; qi < NQ; ++qi)
#pragma unroll
        for (int e = 0; e < 8; ++e) qr[qi][e] = qs[qi * DH + j * 8 + e];

    // ---- phase A: scores; shared positions [0,c): one K row serves all slots ----
    for (int p0 = w * 16; p0 < c; p0 += AT_WARPS * 16) {
        Raw8<T> raw[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = p0 + u * 4 + g;
            if (p < c) raw[u].load(kaddr(0, p));
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = p0 + u * 4 + g;
            float kf[8];
            if (p < c) raw[u].unpack(kf);
            else {
#pragma unroll
                for (int e = 0; e < 8; ++e) kf[e] = 0.f;
            }
#pragma unroll
            for (int qi = 0; qi < NQ; ++qi) {
                float acc = 0.f;
#pragma unroll
                for (int e = 0; e < 8; ++e) acc = fmaf(kf[e], qr[qi][e], acc);
                acc += __shfl_xor_sync(0xffffffffu, acc, 1);
                acc += __shfl_xor_sync(0xffffffffu, acc, 2);
                acc += __shfl_xor_sync(0xffffffffu, acc, 4);
                if (j == 0 && p < c) {
                    // additive mask (1 - m) * -10000 : reference src/bert_seq2seq.py:36-37
                    if (!SELF && a.mask != nullptr && a.mask[(size_t)b * a.S + p] == 0) acc += -10000.0f;
                    sc[qi * npad + p] = acc;
                }
            }
        }
    }
    // divergent tail [c,n) (self-attention only): lane group g serves slot g, its positions dealt to the warps
    float qd[8];
    if (SELF && NQ > 1) {
#pragma unroll
        for (int e = 0; e < 8; ++e) qd[e] = qs[min(g, NQ - 1) * DH + j * 8 + e];
        for (int p0 = c + w * 4; p0 < n; p0 += AT_WARPS * 4) {
            Raw8<T> raw[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int p = p0 + u;
                if (p < n && g < NQ) raw[u].load(kaddr(g, p));
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int p = p0 + u;
                float kf[8];
                if (p < n && g < NQ) raw[u].unpack(kf);
                else {
#pragma unroll
                    for (int e = 0; e < 8; ++e) kf[e] = 0.f;
                }
                float acc = 0.f;
#pragma unroll
                for (int e = 0; e < 8; ++e) acc = fmaf(kf[e], qd[e], acc);
                acc += __shfl_xor_sync(0xffffffffu, acc, 1);
                acc += __shfl_xor_sync(0xffffffffu, acc, 2);
                acc += __shfl_xor_sync(0xffffffffu, acc, 4);
                if (j == 0 && p < n && g < NQ) sc[g * npad + p] = acc;
            }
        }
    }
    __syncthreads();
    // ---- phase B: softmax statistics, one warp per slot ----
    for (int qi = w; qi < NQ; qi += AT_WARPS) {
        float* s = sc + qi * npad;
        float mx = -INFINITY;
        for (int p = lane; p < n; p += 32) mx = fmaxf(mx, s[p]);
        mx = warp_max(mx);
        float sum = 0.f;
        for (int p = lane; p < n; p += 32) {
            float e = expf(s[p] - mx);
            s[p] = e;
            sum += e;
        }
        sum = warp_sum(sum);
        if (lane == 0) inv[qi] = 1.0f / sum;
    }
    __syncthreads();
    // ---- phase C: P.V ----
    float acc[NQ][8];
#pragma unroll
    for (int qi = 0; qi < NQ; ++qi)
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[qi][e] = 0.f;
    for (int p0 = w * 16; p0 < c; p0 += AT_WARPS * 16) {
        Raw8<T> raw[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = p0 + u * 4 + g;
            if (p < c) raw[u].load(kaddr(0, p) + d);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = p0 + u * 4 + g;
            if (p < c) {
                float vf[8];
                raw[u].unpack(vf);
#pragma unroll
                for (int qi = 0; qi < NQ; ++qi) {
                    const float e = sc[qi * npad + p];
#pragma unroll
                    for (int k = 0; k < 8; ++k) acc[qi][k] = fmaf(e, vf[k], acc[qi][k]);
                }
            }
        }
    }
#pragma unroll
    for (int qi = 0; qi < NQ; ++qi)
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float v = acc[qi][k];
            v += __shfl_xor_sync(0xffffffffu, v, 8);
            v += __shfl_xor_sync(0xffffffffu, v, 16);
            if (g == 0) pv[(w * NQ + qi) * DH + j * 8 + k] = v;
        }
    if (SELF && NQ > 1) {
        float ad[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) ad[k] = 0.f;
        for (int p0 = c + w * 4; p0 < n; p0 += AT_WARPS * 4) {
            Raw8<T> raw[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int p = p0 + u;
                if (p < n && g < NQ) raw[u].load(kaddr(g, p) + d);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int p = p0 + u;
                if (p < n && g < NQ) {
                    float vf[8];
                    raw[u].unpack(vf);
                    const float e = sc[g * npad + p];
#pragma unroll
                    for (int k = 0; k < 8; ++k) ad[k] = fmaf(e, vf[k], ad[k]);
                }
            }
        }
        __syncwarp();
        if (g < NQ) {
#pragma unroll
            for (int k = 0; k < 8; ++k) pv[(w * NQ + g) * DH + j * 8 + k] += ad[k];
        }
    }
    __syncthreads();
    for (int i = tid; i < nq * DH; i += AT_THREADS) {
        const int qi = i / DH, k = i % DH;
        float v = 0.f;
#pragma unroll
        for (int ww = 0; ww < AT_WARPS; ++ww) v += pv[(ww * NQ + qi) * DH + k];
        reinterpret_cast<T*>(a.ctx)[(size_t)(row0 + qi) * d + h * DH + k] = from_f32<T>(v * inv[qi]);
    }
}

template <typename T>
static int launch_self(const AttnArgs& a, int H, cudaStream_t s) {
    const int B = a.M / a.beam;
    const int np = (a.T_max + 3) & ~3;
    cudaError_t e;
    if (a.beam == 1) {
        const size_t smem = ((size_t)np + AT_WARPS * DH + DH + 1) * sizeof(float) + a.T_max + 16;
        e = itb_launch(attn_kernel<T, 1, true>, dim3(B, H, 1), dim3(AT_THREADS), smem, s, a);
    } else {
        const size_t smem = ((size_t)NQ_MAX * np + AT_WARPS * NQ_MAX * DH + NQ_MAX * DH + NQ_MAX) * sizeof(float) + (size_t)NQ_MAX * a.T_max + 16;
        e = itb_launch(attn_kernel<T, NQ_MAX, true>, dim3(B, H, (a.beam + NQ_MAX - 1) / NQ_MAX), dim3(AT_THREADS), smem, s, a);
    }
    if (e != cudaSuccess) { itb_set_error("self_attention: launch failed: %s", cudaGetErrorString(e)); return -1; }
    return itb_check_launch("self_attention");
}

template <typename T>
static int launch_cross(const AttnArgs& a, int H, cudaStream_t s) {
    const int B = a.M / a.beam;
    const int np = (a.S + 3) & ~3;
    cudaError_t e;
    if (a.beam == 1) {
        const size_t smem = ((size_t)np + AT_WARPS * DH + DH + 1) * sizeof(float) + 16;
        e = itb_launch(attn_kernel<T, 1, false>, dim3(B, H, 1), dim3(AT_THREADS), smem, s, a);
    } else {
        const size_t smem = ((size_t)NQ_MAX * np + AT_WARPS * NQ_MAX * DH + NQ_MAX * DH + NQ_MAX) * sizeof(float) + 16;
        e = itb_launch(attn_kernel<T, NQ_MAX, false>, dim3(B, H, (a.beam + NQ_MAX - 1) / NQ_MAX), dim3(AT_THREADS), smem, s, a);
    }
    if (e != cudaSuccess) { itb_set_error("cross_attention: launch failed: %s", cudaGetErrorString(e)); return -1; }
    return itb_check_launch("cross_attention");
}

int itb_attn_mma(int dtype, bool self, const void* ctl, const int* sent_done, const int* mem_len, const float* q, int ldq,
                 void* kv, const unsigned char* anc, const unsigned char* mask, void* ctx, int M, int beam, int d, int T_max,
                 int S, int L, int layer, int early_kv, cudaStream_t s);

// 16-bit modes run the tensor-core kernel (attention_mma.cu); fp32 (parity mode) and ITB_ATTN=simt run the CUDA-core one.
static bool use_mma(int dtype) {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("ITB_ATTN");
        v = (e != nullptr && e[0] == 's') ? 0 : 1;
    }
    return v != 0 && dtype != ITB_F32;
}

int itb_cross_attention_impl(int dtype, const itb_ctl* ctl, const int* sent_done, const int* mem_len, const float* q, int ldq,
                             const void* cross_kv, int layer, const unsigned char* src_mask, void* ctx, int M, int beam, int d,
                             int H, int S, int L, int early_kv, void* stream);

extern "C" {

int itb_self_attention(int dtype, const itb_ctl* ctl, const int* sent_done, const float* qkv, void* kv_layer,
                       const unsigned char* anc, void* ctx, int M, int beam, int d, int H, int T_max, void* stream) {
    ITB_REQUIRE(d == H * DH, "self_attention: head_dim must be 64 (d == 64*H)");
    ITB_REQUIRE(T_max <= 2048, "self_attention: T_max too large for the shared-memory score buffer");
    ITB_REQUIRE(M % beam == 0, "self_attention: M must be a multiple of the beam width");
    if (use_mma(dtype))
        return itb_attn_mma(dtype, true, ctl, sent_done, nullptr, qkv, 3 * d, kv_layer, anc, nullptr, ctx, M, beam, d, T_max, 0, 0, 0, 0,
                            (cudaStream_t)stream);
    AttnArgs a = {};
    a.ctl = reinterpret_cast<const ItbCtl*>(ctl);
    a.sent_done = sent_done;
    a.q = qkv; a.ldq = 3 * d; a.kv = kv_layer; a.anc = anc; a.ctx = ctx;
    a.M = M; a.beam = beam; a.d = d; a.T_max = T_max;
    if (dtype == ITB_BF16) return launch_self<__nv_bfloat16>(a, H, (cudaStream_t)stream);
    if (dtype == ITB_F16) return launch_self<__half>(a, H, (cudaStream_t)stream);
    return launch_self<float>(a, H, (cudaStream_t)stream);
}

int itb_cross_attention(int dtype, const itb_ctl* ctl, const int* sent_done, const int* mem_len, const float* q, int ldq,
                        const void* cross_kv, int layer, const unsigned char* src_mask, void* ctx, int M, int beam, int d,
                        int H, int S, int L, void* stream) {
    return itb_cross_attention_impl(dtype, ctl, sent_done, mem_len, q, ldq, cross_kv, layer, src_mask, ctx, M, beam, d, H, S, L, 0, stream);
}

}  // extern "C"

// early_kv != 0: the caller guarantees that cross_kv and src_mask were complete before the last FULL stream dependency
// (the decoder: cross-KV GEMM -> init_state (plain launch) -> steps), so the kernel may fetch them ahead of its
// programmatic-dependency wait.  The public entry point cannot know that and passes 0.
int itb_cross_attention_impl(int dtype, const itb_ctl* ctl, const int* sent_done, const int* mem_len, const float* q, int ldq,
                             const void* cross_kv, int layer, const unsigned char* src_mask, void* ctx, int M, int beam, int d,
                             int H, int S, int L, int early_kv, void* stream) {
    ITB_REQUIRE(d == H * DH, "cross_attention: head_dim must be 64 (d == 64*H)");
    ITB_REQUIRE(S <= 2048, "cross_attention: S too large for the shared-memory score buffer");
    ITB_REQUIRE(M % beam == 0, "cross_attention: M must be a multiple of the beam width");
    if (use_mma(dtype))
        return itb_attn_mma(dtype, false, ctl, sent_done, mem_len, q, ldq, const_cast<void*>(cross_kv), nullptr, src_mask, ctx, M, beam, d,
                            0, S, L, layer, early_kv, (cudaStream_t)stream);
    AttnArgs a = {};
    a.ctl = reinterpret_cast<const ItbCtl*>(ctl);
    a.sent_done = sent_done; a.mem_len = mem_len;
    a.q = q; a.ldq = ldq; a.kv = const_cast<void*>(cross_kv); a.mask = src_mask; a.ctx = ctx;
    a.M = M; a.beam = beam; a.d = d; a.S = S; a.L = L; a.layer = layer;
    if (dtype == ITB_BF16) return launch_cross<__nv_bfloat16>(a, H, (cudaStream_t)stream);
    if (dtype == ITB_F16) return launch_cross<__half>(a, H, (cudaStream_t)stream);
    return launch_cross<float>(a, H, (cudaStream_t)stream);
}


