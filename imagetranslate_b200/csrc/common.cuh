// Shared device helpers for the ImageTranslate B200 decode path (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

#define ITB_WARP 32

// Device-resident control block shared by every kernel of a decode call.  All step kernels
// read `step`/`done` from here so that ONE captured CUDA graph can be replayed for every
// step, and so that steps enqueued past termination are exact no-ops
// (termination rule: /root/reference/src/seq_gen.py:134-136).
struct ItbCtl {
    int step;        // current 1-based step index i of seq_gen.py:131 (the token consumed is at position i-1)
    int done;        // 1 once every row of the batch holds an EOS (checked before each step, like :134)
    int steps_run;   // number of steps actually executed
    int cur;         // ping-pong parity of the token / ancestor tables
    int fin_rows;    // rows (over the whole batch) whose prefix holds an EOS
    int sent_done;   // sentences that finished the merge phase of the current step
    int max_len;     // loop bound of :131 (steps run for i in [1, max_len))
    int rows_active; // rows [rows_active, M) belong to groups whose search has ended (groups are laid out longest loop bound
                     // first, so finished rows form a suffix): the dense kernels skip them.  <= 0: unset, every row is live
};

// Programmatic dependent launch (PDL).  Every kernel of the step chain is launched with the programmatic-stream-
// serialization attribute: it may START while its predecessor is still running, executes only work that does not
// depend on earlier kernels (barrier / TMEM set-up, weight-tile TMA loads), and calls pdl_wait() before touching
// anything an earlier kernel wrote (or writing anything an earlier kernel reads).  pdl_wait() returns once the
// predecessor grid has completed and flushed; pdl_trigger() lets the successor be scheduled.
// Per-group (= per reference batch) control block of a super-batch; mirrors itb_grp in include/itb.h.
struct ItbGrp {
    int done;        // the group's loop has ended (every row holds EOS before a step, or step reached max_len)
    int steps_run;   // steps executed for this group
    int max_len;     // loop bound of seq_gen.py:131 for this group
    int fin_rows;    // accumulator: rows of the group whose NEW prefix holds an EOS
    int n_rows;      // rows (sentences * beam) of the group
    int final_cur;   // ping-pong parity holding the group's final state
    int pad0_, pad1_;
};

// rows the dense kernels still have to compute (see ItbCtl::rows_active)
__device__ __forceinline__ int itb_rows_active(const ItbCtl* ctl) {
    if (ctl == nullptr) return 0x7fffffff;
    const int ra = ctl->rows_active;
    return ra > 0 ? ra : 0x7fffffff;
}

#ifndef ITB_TIMELINE
#define ITB_TL(n) do { } while (0)
#define ITB_TL_ANY(n) do { } while (0)
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#else
// Diagnostic build (csrc/timeline_all.cu, never the shipped library): thread 0 of block 0 of every kernel stamps
// %globaltimer when it enters the kernel (at pdl_trigger) and when its programmatic dependency resolves (after pdl_wait), so
// that the chain of a real search step can be laid out launch by launch.  Entry = (source tag, time); tag = file * 100000 +
// line (+ 50000 for the dependency stamp).
#define ITB_TL_CAP 16384
__device__ unsigned long long g_itb_tl[1 + 2 * ITB_TL_CAP];
__device__ __forceinline__ void itb_tl_stamp(int tag) {
    if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        const unsigned long long i = atomicAdd(&g_itb_tl[0], 1ull);
        if (i < ITB_TL_CAP) { g_itb_tl[1 + 2 * i] = (unsigned long long)tag; g_itb_tl[2 + 2 * i] = t; }
    }
}
__device__ __forceinline__ void pdl_wait_impl() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger_impl() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// extra stamps inside a kernel (diagnostic build only): ITB_TL(n) by block 0, ITB_TL_ANY(n) by thread 0 of whichever block runs it
#define ITB_TL(n) itb_tl_stamp(ITB_TL_FILE * 100000 + 90000 + (n))
#define ITB_TL_ANY(n) do { if (threadIdx.x == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); \
        const unsigned long long i_ = atomicAdd(&g_itb_tl[0], 1ull); \
        if (i_ < ITB_TL_CAP) { g_itb_tl[1 + 2 * i_] = (unsigned long long)(ITB_TL_FILE * 100000 + 90000 + (n)); g_itb_tl[2 + 2 * i_] = t_; } } } while (0)
#define pdl_wait() do { pdl_wait_impl(); itb_tl_stamp(ITB_TL_FILE * 100000 + 50000 + __LINE__); } while (0)
#define pdl_trigger() do { itb_tl_stamp(ITB_TL_FILE * 100000 + __LINE__); pdl_trigger_impl(); } while (0)
#endif

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide sum for blockDim.x <= 1024; `red` is >= 32 floats of shared memory.
__device__ __forceinline__ float block_sum(float v, float* red) {
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    float t = (threadIdx.x < nw) ? red[threadIdx.x] : 0.f;
    if (w == 0) {
        t = warp_sum(t);
        if (lane == 0) red[0] = t;
    }
    __syncthreads();
    return red[0];
}

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// exact (erf) GELU, as HF "gelu" (transformers 2.9.0 BertIntermediate)
__device__ __forceinline__ float gelu_erf(float x) { return x * 0.5f * (1.0f + erff(x * 0.70710678118654752440f)); }

// erf GELU for the 16-bit tensor-core epilogues, one MUFU and no branch: erf(z) = 1 - (1 + a1 z + ... + a6 z^6)^-16
// (Abramowitz-Stegun 7.1.28, |error| <= 3e-7; GELU abs error < 1e-6, far below the bf16/fp16 rounding of the result).
// The fp32 parity path keeps erff (gelu_erf).
__device__ __forceinline__ float gelu_erf_fast(float x) {
    const float z = fabsf(x) * 0.70710678118654752440f;
    float p = fmaf(z, 0.0000430638f, 0.0002765672f);
    p = fmaf(z, p, 0.0001520143f);
    p = fmaf(z, p, 0.0092705272f);
    p = fmaf(z, p, 0.0422820123f);
    p = fmaf(z, p, 0.0705230784f);
    p = fmaf(z, p, 1.0f);
    p *= p; p *= p; p *= p; p *= p;                       // ^16 (overflows to +inf for z > ~17: erf = 1, as it should)
    const float e = 1.0f - __fdividef(1.0f, p);
    const float hx = 0.5f * x;
    return fmaf(hx, copysignf(e, x), hx);
}
