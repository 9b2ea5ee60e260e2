// CUDA-core GEMM  out[M][N] = epi(A[M][K] . W[N][K]^T + bias)  -- the fp32 parity path of the decoder
// (ITB_F32) and a bring-up cross-check for the tcgen05 kernel.  The bf16 production path is gemm_tc.cu.
// 64x64x32 tiles, 256 threads, 4x4 register micro-tiles, smem tiles stored k-major (conflict-free float4 reads).
#include "common.cuh"
#include "launch.h"

#define BM 64
#define BN 64
#define BK 32

template <typename T> struct Ld8;   // loads 8 consecutive K-elements as floats
template <> struct Ld8<float> {
    static __device__ __forceinline__ void ld(const float* p, float* o) {
        float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
        o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
    }
};
template <> struct Ld8<__nv_bfloat16> {
    static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float* o) {
        uint4 u = *reinterpret_cast<const uint4*>(p);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
        for (int e = 0; e < 4; ++e) { float2 f = __bfloat1622float2(h[e]); o[2 * e] = f.x; o[2 * e + 1] = f.y; }
    }
};

template <> struct Ld8<__half> {
    static __device__ __forceinline__ void ld(const __half* p, float* o) {
        uint4 u = *reinterpret_cast<const uint4*>(p);
        const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
        for (int e = 0; e < 4; ++e) { float2 f = __half22float2(h[e]); o[2 * e] = f.x; o[2 * e + 1] = f.y; }
    }
};

template <typename T, typename TW, int EPI>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const ItbCtl* __restrict__ ctl, const T* __restrict__ A,
                                                        const TW* __restrict__ W, const float* __restrict__ bias,
                                                        void* __restrict__ out, int M, int N, int K, int ldo) {
    __shared__ float As[BK][BM + 4];
    __shared__ float Ws[BK][BN + 4];
    if (ctl != nullptr && ctl->done) return;
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int tx = tid & 15, ty = tid >> 4;          // 16 x 16 threads, 4x4 outputs each
    const int lr = tid >> 2, lk = (tid & 3) * 8;     // loader: row 0..63, k offset 0,8,16,24
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (int k0 = 0; k0 < K; k0 += BK) {
        float av[8], wv[8];
        if (m0 + lr < M) Ld8<T>::ld(A + (size_t)(m0 + lr) * K + k0 + lk, av);
        else {
#pragma unroll
            for (int e = 0; e < 8; ++e) av[e] = 0.f;
        }
        if (n0 + lr < N) Ld8<TW>::ld(W + (size_t)(n0 + lr) * K + k0 + lk, wv);
        else {
#pragma unroll
            for (int e = 0; e < 8; ++e) wv[e] = 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int e = 0; e < 8; ++e) { As[lk + e][lr] = av[e]; Ws[lk + e][lr] = wv[e]; }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float4 a4 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            float4 w4 = *reinterpret_cast<const float4*>(&Ws[k][tx * 4]);
            float a[4] = {a4.x, a4.y, a4.z, a4.w}, wv4[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], wv4[j], acc[i][j]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tx * 4 + j;
            if (n >= N) continue;
            float v = acc[i][j] + (bias != nullptr ? bias[n] : 0.f);
            if (EPI == ITB_EPI_BIAS_F32) reinterpret_cast<float*>(out)[(size_t)m * ldo + n] = v;
            else {
                if (EPI == ITB_EPI_BIAS_GELU_ACT) v = gelu_erf(v);
                reinterpret_cast<T*>(out)[(size_t)m * ldo + n] = from_f32<T>(v);
            }
        }
    }
}

template <typename T, typename TW>
static int launch_simt(int epi, const ItbCtl* ctl, const void* A, const void* W, const float* bias, void* out, int M,
                       int N, int K, int ldo, cudaStream_t s) {
    dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM);
    const T* a = (const T*)A; const TW* w = (const TW*)W;
    if (epi == ITB_EPI_BIAS_F32) gemm_simt_kernel<T, TW, ITB_EPI_BIAS_F32><<<grid, 256, 0, s>>>(ctl, a, w, bias, out, M, N, K, ldo);
    else if (epi == ITB_EPI_BIAS_GELU_ACT) gemm_simt_kernel<T, TW, ITB_EPI_BIAS_GELU_ACT><<<grid, 256, 0, s>>>(ctl, a, w, bias, out, M, N, K, ldo);
    else gemm_simt_kernel<T, TW, ITB_EPI_BIAS_ACT><<<grid, 256, 0, s>>>(ctl, a, w, bias, out, M, N, K, ldo);
    return itb_check_launch("gemm_simt");
}

int itb_gemm_simt(int dtype, int epi, const void* ctl, const void* A, const void* W, const float* bias, void* out,
                  int M, int N, int K, int ldo, cudaStream_t s) {
    ITB_REQUIRE(K % BK == 0, "linear: K must be a multiple of 32");
    ITB_REQUIRE(M > 0 && N > 0, "linear: empty problem");
    const ItbCtl* c = reinterpret_cast<const ItbCtl*>(ctl);
    if (dtype == ITB_BF16) return launch_simt<__nv_bfloat16, __nv_bfloat16>(epi, c, A, W, bias, out, M, N, K, ldo, s);
    if (dtype == ITB_F16) return launch_simt<__half, __half>(epi, c, A, W, bias, out, M, N, K, ldo, s);
    return launch_simt<float, float>(epi, c, A, W, bias, out, M, N, K, ldo, s);
}
