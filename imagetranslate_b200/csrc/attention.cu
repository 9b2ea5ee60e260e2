// Single-query attention for the incremental decoder step (head_dim fixed at 64 = 768/12,
// reference src/lm_config.py:8,13; HF BertSelfAttention).
//
// One CTA (4 warps) per (row, head) for self-attention and per (sentence, head, group of <=4 beam slots) for
// cross-attention -- the beam slots of a sentence share the memory K/V, so each K/V byte is fetched once per group.
// Positions are split over the warps in runs of 16; inside a warp 8 lanes x 16 B cover one 64-wide head row and
// 4 positions are processed side by side, 4 such loads in flight per lane (64 positions per CTA iteration):
//   phase A  scores  q.k/8 (+ additive -10000 source mask) -> shared memory
//   phase B  softmax statistics in fp32: exp(x - max) / sum, as torch.softmax
//   phase C  sum_p e_p * V[p], reduced over lanes, warps, then normalised and stored in the activation type.
// Self-attention appends the row's new K,V to the cache first and reads the cache through the ancestor table
// written by the beam kernel, so a beam reorder never copies K/V.  HBM-bound: 2*d*sizeof(T) bytes per
// (row, position, layer) for self-attention, per (sentence, position, layer) for cross-attention.
#include "common.cuh"
#include "launch.h"

#define DH 64
#define AT_WARPS 4
#define AT_THREADS (AT_WARPS * 32)
#define NQ_MAX 4

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

// NQ query rows share the K/V rows addressed by `kaddr(p)` (V row = K row + d).
template <typename T, int NQ, bool SELF>
__global__ void __launch_bounds__(AT_THREADS) attn_kernel(AttnArgs a) {
    extern __shared__ float sm[];
    pdl_trigger();
    pdl_wait();
    const ItbCtl* ctl = a.ctl;
    if (ctl != nullptr && ctl->done) return;
    if (a.sent_done != nullptr && a.sent_done[SELF ? blockIdx.x / a.beam : blockIdx.x] != 0) return;
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, g = lane >> 3, j = lane & 7;
    const int h = blockIdx.y, d = a.d;
    int n, row0, nq;
    const T* kbase;
    size_t pos_stride;
    const unsigned char* arow = nullptr;
    if (SELF) {
        const int r = blockIdx.x;
        row0 = r; nq = 1;
        const int pc = ctl->step - 1;                    // position of the new token; attends to 0..pc
        n = pc + 1;
        kbase = reinterpret_cast<const T*>(a.kv) + h * DH;
        pos_stride = (size_t)2 * d;
        arow = a.anc + ((size_t)ctl->cur * a.M + r) * a.T_max;
    } else {
        const int b = blockIdx.x, q0 = blockIdx.z * NQ;
        row0 = b * a.beam + q0; nq = min(NQ, a.beam - q0);
        n = (a.mem_len != nullptr) ? a.mem_len[b] : a.S;
        kbase = reinterpret_cast<const T*>(a.kv) + ((size_t)b * a.S * a.L + a.layer) * 2 * d + h * DH;
        pos_stride = (size_t)a.L * 2 * d;
    }
    const int npad = (n + 3) & ~3;
    float* sc = sm;                                      // [NQ][npad]
    float* pv = sm + NQ * npad;                          // [AT_WARPS][NQ][DH]
    float* inv = pv + AT_WARPS * NQ * DH;                // [NQ]
    unsigned char* s_anc = reinterpret_cast<unsigned char*>(inv + NQ);     // SELF: [n] physical beam slot per position

    if (SELF) {
        // append this row's K,V (rounded to the cache type) at position pc, then treat it like any cached position
        const int r = row0, pc = n - 1;
        T* knew = reinterpret_cast<T*>(a.kv) + (((size_t)r * a.T_max + pc) * 2) * d + h * DH;
        const float* qrow = a.q + (size_t)r * a.ldq + h * DH;
        if (tid < DH) knew[tid] = from_f32<T>(qrow[d + tid]);
        else knew[d + tid - DH] = from_f32<T>(qrow[2 * d + tid - DH]);
        for (int p = tid; p < pc; p += AT_THREADS) s_anc[p] = arow[p];
        if (tid == 0) s_anc[pc] = (unsigned char)(r % a.beam);
        __syncthreads();
    }
    const int base = SELF ? (row0 / a.beam) * a.beam : 0;
    auto kaddr = [&](int p) -> const T* {
        if (SELF) return kbase + ((size_t)(base + s_anc[p]) * a.T_max + p) * pos_stride + j * 8;
        return kbase + (size_t)p * pos_stride + j * 8;
    };

    // queries: this lane's 8 dims, pre-scaled by 1/sqrt(64) (exact: power of two)
    float qr[NQ][8];
#pragma unroll
    for (int qi = 0; qi < NQ; ++qi) {
        const int r = row0 + min(qi, nq - 1);
        const float4* q4 = reinterpret_cast<const float4*>(a.q + (size_t)r * a.ldq + h * DH + j * 8);
        float4 x = q4[0], y = q4[1];
        qr[qi][0] = x.x * 0.125f; qr[qi][1] = x.y * 0.125f; qr[qi][2] = x.z * 0.125f; qr[qi][3] = x.w * 0.125f;
        qr[qi][4] = y.x * 0.125f; qr[qi][5] = y.y * 0.125f; qr[qi][6] = y.z * 0.125f; qr[qi][7] = y.w * 0.125f;
    }

    // ---- phase A: scores ----
    for (int p0 = w * 16; p0 < n; p0 += AT_WARPS * 16) {
        Raw8<T> raw[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = p0 + u * 4 + g;
            if (p < n) raw[u].load(kaddr(p));
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = p0 + u * 4 + g;
            float kf[8];
            if (p < n) raw[u].unpack(kf);
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
                if (j == 0 && p < n) {
                    // additive mask (1 - m) * -10000 : reference src/bert_seq2seq.py:36-37
                    if (!SELF && a.mask != nullptr && a.mask[(size_t)blockIdx.x * a.S + p] == 0) acc += -10000.0f;
                    sc[qi * npad + p] = acc;
                }
            }
        }
    }
    __syncthreads();
    // ---- phase B: softmax statistics, one warp per query ----
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
    for (int p0 = w * 16; p0 < n; p0 += AT_WARPS * 16) {
        Raw8<T> raw[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = p0 + u * 4 + g;
            if (p < n) raw[u].load(kaddr(p) + d);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = p0 + u * 4 + g;
            if (p < n) {
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
    const size_t smem = ((size_t)((a.T_max + 3) & ~3) + AT_WARPS * DH + 1) * sizeof(float) + a.T_max + 16;
    cudaError_t e = itb_launch(attn_kernel<T, 1, true>, dim3(a.M, H), dim3(AT_THREADS), smem, s, a);
    if (e != cudaSuccess) { itb_set_error("self_attention: launch failed: %s", cudaGetErrorString(e)); return -1; }
    return itb_check_launch("self_attention");
}

template <typename T>
static int launch_cross(const AttnArgs& a, int H, cudaStream_t s) {
    const size_t smem = ((size_t)NQ_MAX * ((a.S + 3) & ~3) + AT_WARPS * NQ_MAX * DH + NQ_MAX) * sizeof(float) + 16;
    const int B = a.M / a.beam;
    cudaError_t e;
    if (a.beam == 1) e = itb_launch(attn_kernel<T, 1, false>, dim3(B, H, 1), dim3(AT_THREADS), smem, s, a);
    else if (a.beam == 2) e = itb_launch(attn_kernel<T, 2, false>, dim3(B, H, 1), dim3(AT_THREADS), smem, s, a);
    else e = itb_launch(attn_kernel<T, NQ_MAX, false>, dim3(B, H, (a.beam + NQ_MAX - 1) / NQ_MAX), dim3(AT_THREADS), smem, s, a);
    if (e != cudaSuccess) { itb_set_error("cross_attention: launch failed: %s", cudaGetErrorString(e)); return -1; }
    return itb_check_launch("cross_attention");
}

extern "C" {

int itb_self_attention(int dtype, const itb_ctl* ctl, const int* sent_done, const float* qkv, void* kv_layer,
                       const unsigned char* anc, void* ctx, int M, int beam, int d, int H, int T_max, void* stream) {
    ITB_REQUIRE(d == H * DH, "self_attention: head_dim must be 64 (d == 64*H)");
    ITB_REQUIRE(T_max <= 2048, "self_attention: T_max too large for the shared-memory score buffer");
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
    ITB_REQUIRE(d == H * DH, "cross_attention: head_dim must be 64 (d == 64*H)");
    ITB_REQUIRE(S <= 2048, "cross_attention: S too large for the shared-memory score buffer");
    ITB_REQUIRE(M % beam == 0, "cross_attention: M must be a multiple of the beam width");
    AttnArgs a = {};
    a.ctl = reinterpret_cast<const ItbCtl*>(ctl);
    a.sent_done = sent_done; a.mem_len = mem_len;
    a.q = q; a.ldq = ldq; a.kv = const_cast<void*>(cross_kv); a.mask = src_mask; a.ctx = ctx;
    a.M = M; a.beam = beam; a.d = d; a.S = S; a.L = L; a.layer = layer;
    if (dtype == ITB_BF16) return launch_cross<__nv_bfloat16>(a, H, (cudaStream_t)stream);
    if (dtype == ITB_F16) return launch_cross<__half>(a, H, (cudaStream_t)stream);
    return launch_cross<float>(a, H, (cudaStream_t)stream);
}

}  // extern "C"
