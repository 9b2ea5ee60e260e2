// Single-query attention for the incremental decoder step (head_dim fixed at 64 = 768/12,
// reference src/lm_config.py:8,13).  One warp per (row, head):
//   phase A  lane-per-position  q.k  (each lane streams one full 128/256-byte K row)
//   phase B  warp softmax statistics in fp32 (exp(x-max)/sum, as torch.softmax)
//   phase C  lane-per-dim       sum_p e_p * V[p]  (coalesced 128/256-byte V rows)
// Self-attention reads the KV cache through the ancestor table written by the beam kernel, so a
// beam reorder never copies K/V.  HBM-bound: 2*d*sizeof(T) bytes per (row, position, layer).
#include "common.cuh"
#include "launch.h"

#define DH 64

template <typename T> struct RowLoad;
template <> struct RowLoad<float> {
    static __device__ __forceinline__ float dot(const float* k, const float* q) {
        const float4* k4 = reinterpret_cast<const float4*>(k);
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < DH / 4; ++j) {
            float4 v = k4[j];
            acc = fmaf(v.x, q[4 * j + 0], acc);
            acc = fmaf(v.y, q[4 * j + 1], acc);
            acc = fmaf(v.z, q[4 * j + 2], acc);
            acc = fmaf(v.w, q[4 * j + 3], acc);
        }
        return acc;
    }
    static __device__ __forceinline__ float2 pair(const float* v, int lane) {
        return reinterpret_cast<const float2*>(v)[lane];
    }
};
template <> struct RowLoad<__nv_bfloat16> {
    static __device__ __forceinline__ float dot(const __nv_bfloat16* k, const float* q) {
        const uint4* k4 = reinterpret_cast<const uint4*>(k);
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < DH / 8; ++j) {
            uint4 u = k4[j];
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                float2 f = __bfloat1622float2(h[e]);
                acc = fmaf(f.x, q[8 * j + 2 * e + 0], acc);
                acc = fmaf(f.y, q[8 * j + 2 * e + 1], acc);
            }
        }
        return acc;
    }
    static __device__ __forceinline__ float2 pair(const __nv_bfloat16* v, int lane) {
        return __bfloat1622float2(reinterpret_cast<const __nv_bfloat162*>(v)[lane]);
    }
};

// Shared tail: softmax over sc[0..n) (+ optional extra current-position score) and P.V accumulation.
template <> struct RowLoad<__half> {
    static __device__ __forceinline__ float dot(const __half* k, const float* q) {
        const uint4* k4 = reinterpret_cast<const uint4*>(k);
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < DH / 8; ++j) {
            uint4 u = k4[j];
            const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                float2 f = __half22float2(h[e]);
                acc = fmaf(f.x, q[8 * j + 2 * e + 0], acc);
                acc = fmaf(f.y, q[8 * j + 2 * e + 1], acc);
            }
        }
        return acc;
    }
    static __device__ __forceinline__ float2 pair(const __half* v, int lane) {
        return __half22float2(reinterpret_cast<const __half2*>(v)[lane]);
    }
};

template <typename T, typename VPtr>
__device__ __forceinline__ void softmax_pv(float* sc, int n, bool has_cur, float s_cur, float2 v_cur, VPtr vptr,
                                           int lane, float2& out) {
    float mx = has_cur ? s_cur : -INFINITY;
    for (int p = lane; p < n; p += 32) mx = fmaxf(mx, sc[p]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int p = lane; p < n; p += 32) {
        float e = expf(sc[p] - mx);
        sc[p] = e;
        sum += e;
    }
    sum = warp_sum(sum);
    float e_cur = 0.f;
    if (has_cur) { e_cur = expf(s_cur - mx); sum += e_cur; }
    __syncwarp();
    float ax = 0.f, ay = 0.f, bx = 0.f, by = 0.f;
    int p = 0;
    for (; p + 1 < n; p += 2) {
        float2 v0 = RowLoad<T>::pair(vptr(p), lane);
        float2 v1 = RowLoad<T>::pair(vptr(p + 1), lane);
        float e0 = sc[p], e1 = sc[p + 1];
        ax = fmaf(e0, v0.x, ax); ay = fmaf(e0, v0.y, ay);
        bx = fmaf(e1, v1.x, bx); by = fmaf(e1, v1.y, by);
    }
    if (p < n) {
        float2 v0 = RowLoad<T>::pair(vptr(p), lane);
        float e0 = sc[p];
        ax = fmaf(e0, v0.x, ax); ay = fmaf(e0, v0.y, ay);
    }
    if (has_cur) { ax = fmaf(e_cur, v_cur.x, ax); ay = fmaf(e_cur, v_cur.y, ay); }
    float inv = 1.0f / sum;
    out.x = (ax + bx) * inv;
    out.y = (ay + by) * inv;
}

template <typename T, int HPC>
__global__ void __launch_bounds__(HPC * 32) self_attn_kernel(const ItbCtl* __restrict__ ctl, const float* __restrict__ qkv,
                                                             T* kv, const unsigned char* __restrict__ anc,
                                                             T* __restrict__ ctx, int M, int beam, int d, int T_max) {
    extern __shared__ float sm[];
    if (ctl->done) return;
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r = blockIdx.x, h = blockIdx.y * HPC + w;
    const int pc = ctl->step - 1;                    // position of the new token; attends to 0..pc
    float* qs = sm + w * DH;
    float* sc = sm + HPC * DH + w * T_max;
    const float* qrow = qkv + (size_t)r * 3 * d + h * DH;
    // new K,V rounded to the cache type, appended at (row r, position pc)
    float2 q2 = reinterpret_cast<const float2*>(qrow)[lane];
    float2 k2 = reinterpret_cast<const float2*>(qrow + d)[lane];
    float2 v2 = reinterpret_cast<const float2*>(qrow + 2 * d)[lane];
    T kx = from_f32<T>(k2.x), ky = from_f32<T>(k2.y), vx = from_f32<T>(v2.x), vy = from_f32<T>(v2.y);
    T* knew = kv + (((size_t)r * T_max + pc) * 2) * d + h * DH;
    knew[2 * lane] = kx; knew[2 * lane + 1] = ky;
    knew[d + 2 * lane] = vx; knew[d + 2 * lane + 1] = vy;
    qs[2 * lane] = q2.x; qs[2 * lane + 1] = q2.y;
    float s_cur = warp_sum(q2.x * to_f32<T>(kx) + q2.y * to_f32<T>(ky)) * 0.125f;
    __syncwarp();
    const unsigned char* arow = anc + ((size_t)ctl->cur * M + r) * T_max;
    const int base = (r / beam) * beam;
    for (int p = lane; p < pc; p += 32) {
        const T* k = kv + (((size_t)(base + arow[p]) * T_max + p) * 2) * d + h * DH;
        sc[p] = RowLoad<T>::dot(k, qs) * 0.125f;
    }
    __syncwarp();
    float2 o;
    auto vptr = [&](int p) { return kv + (((size_t)(base + arow[p]) * T_max + p) * 2 + 1) * d + h * DH; };
    softmax_pv<T>(sc, pc, true, s_cur, make_float2(to_f32<T>(vx), to_f32<T>(vy)), vptr, lane, o);
    T* orow = ctx + (size_t)r * d + h * DH;
    orow[2 * lane] = from_f32<T>(o.x);
    orow[2 * lane + 1] = from_f32<T>(o.y);
}

template <typename T, int HPC>
__global__ void __launch_bounds__(HPC * 32) cross_attn_kernel(const ItbCtl* __restrict__ ctl, const float* __restrict__ q,
                                                              int ldq, const T* __restrict__ ckv, int layer,
                                                              const unsigned char* __restrict__ mask,
                                                              T* __restrict__ ctx, int M, int beam, int d, int S, int L) {
    extern __shared__ float sm[];
    if (ctl != nullptr && ctl->done) return;
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r = blockIdx.x, h = blockIdx.y * HPC + w;
    const int b = r / beam;
    float* qs = sm + w * DH;
    float* sc = sm + HPC * DH + w * S;
    float2 q2 = reinterpret_cast<const float2*>(q + (size_t)r * ldq + h * DH)[lane];
    qs[2 * lane] = q2.x; qs[2 * lane + 1] = q2.y;
    __syncwarp();
    const size_t pos_stride = (size_t)L * 2 * d;
    const T* kbase = ckv + ((size_t)b * S * L + layer) * 2 * d + h * DH;
    for (int p = lane; p < S; p += 32) {
        float s = RowLoad<T>::dot(kbase + p * pos_stride, qs) * 0.125f;
        // additive mask (1 - m) * -10000 : reference src/bert_seq2seq.py:36-37
        if (mask != nullptr && mask[(size_t)b * S + p] == 0) s += -10000.0f;
        sc[p] = s;
    }
    __syncwarp();
    float2 o;
    auto vptr = [&](int p) { return kbase + p * pos_stride + d; };
    softmax_pv<T>(sc, S, false, 0.f, make_float2(0.f, 0.f), vptr, lane, o);
    T* orow = ctx + (size_t)r * d + h * DH;
    orow[2 * lane] = from_f32<T>(o.x);
    orow[2 * lane + 1] = from_f32<T>(o.y);
}

static int pick_hpc(int H) { return (H % 4 == 0) ? 4 : ((H % 2 == 0) ? 2 : 1); }

template <typename T>
static int launch_self(const ItbCtl* ctl, const float* qkv, void* kv, const unsigned char* anc, void* ctx, int M,
                       int beam, int d, int H, int T_max, cudaStream_t s) {
    int hpc = pick_hpc(H);
    size_t smem = (size_t)hpc * (DH + T_max) * sizeof(float);
    dim3 grid(M, H / hpc);
    if (hpc == 4) self_attn_kernel<T, 4><<<grid, 128, smem, s>>>(ctl, qkv, (T*)kv, anc, (T*)ctx, M, beam, d, T_max);
    else if (hpc == 2) self_attn_kernel<T, 2><<<grid, 64, smem, s>>>(ctl, qkv, (T*)kv, anc, (T*)ctx, M, beam, d, T_max);
    else self_attn_kernel<T, 1><<<grid, 32, smem, s>>>(ctl, qkv, (T*)kv, anc, (T*)ctx, M, beam, d, T_max);
    return itb_check_launch("self_attention");
}

template <typename T>
static int launch_cross(const ItbCtl* ctl, const float* q, int ldq, const void* ckv, int layer,
                        const unsigned char* mask, void* ctx, int M, int beam, int d, int H, int S, int L,
                        cudaStream_t s) {
    int hpc = pick_hpc(H);
    size_t smem = (size_t)hpc * (DH + S) * sizeof(float);
    dim3 grid(M, H / hpc);
    if (hpc == 4) cross_attn_kernel<T, 4><<<grid, 128, smem, s>>>(ctl, q, ldq, (const T*)ckv, layer, mask, (T*)ctx, M, beam, d, S, L);
    else if (hpc == 2) cross_attn_kernel<T, 2><<<grid, 64, smem, s>>>(ctl, q, ldq, (const T*)ckv, layer, mask, (T*)ctx, M, beam, d, S, L);
    else cross_attn_kernel<T, 1><<<grid, 32, smem, s>>>(ctl, q, ldq, (const T*)ckv, layer, mask, (T*)ctx, M, beam, d, S, L);
    return itb_check_launch("cross_attention");
}

extern "C" {

int itb_self_attention(int dtype, const itb_ctl* ctl, const float* qkv, void* kv_layer, const unsigned char* anc,
                       void* ctx, int M, int beam, int d, int H, int T_max, void* stream) {
    ITB_REQUIRE(d == H * DH, "self_attention: head_dim must be 64 (d == 64*H)");
    ITB_REQUIRE(T_max <= 2048, "self_attention: T_max too large for the shared-memory score buffer");
    const ItbCtl* c = reinterpret_cast<const ItbCtl*>(ctl);
    if (dtype == ITB_BF16) return launch_self<__nv_bfloat16>(c, qkv, kv_layer, anc, ctx, M, beam, d, H, T_max, (cudaStream_t)stream);
    if (dtype == ITB_F16) return launch_self<__half>(c, qkv, kv_layer, anc, ctx, M, beam, d, H, T_max, (cudaStream_t)stream);
    return launch_self<float>(c, qkv, kv_layer, anc, ctx, M, beam, d, H, T_max, (cudaStream_t)stream);
}

int itb_cross_attention(int dtype, const itb_ctl* ctl, const float* q, int ldq, const void* cross_kv, int layer,
                        const unsigned char* src_mask, void* ctx, int M, int beam, int d, int H, int S, int L,
                        void* stream) {
    ITB_REQUIRE(d == H * DH, "cross_attention: head_dim must be 64 (d == 64*H)");
    ITB_REQUIRE(S <= 2048, "cross_attention: S too large for the shared-memory score buffer");
    const ItbCtl* c = reinterpret_cast<const ItbCtl*>(ctl);
    if (dtype == ITB_BF16) return launch_cross<__nv_bfloat16>(c, q, ldq, cross_kv, layer, src_mask, ctx, M, beam, d, H, S, L, (cudaStream_t)stream);
    if (dtype == ITB_F16) return launch_cross<__half>(c, q, ldq, cross_kv, layer, src_mask, ctx, M, beam, d, H, S, L, (cudaStream_t)stream);
    return launch_cross<float>(c, q, ldq, cross_kv, layer, src_mask, ctx, M, beam, d, H, S, L, (cudaStream_t)stream);
}

}  // extern "C"
