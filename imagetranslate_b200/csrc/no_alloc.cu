// itb_linear_add_ln_static: x = LayerNorm(A . W^T + bias + x) with the GEMM's split-K partial outputs parked in a
// file-scope device array instead of a caller-allocated `y` buffer.  For callers that cannot allocate device memory
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The arithmetic is the decode path's own: the projection GEMM with its planned
// split-K factor (itb_gemm_tc / itb_gemm_simt) followed by itb_add_ln over the partials (HF BertSelfOutput / BertOutput).
//
// The array is ONE per GPU per loaded library and lives until the process exits.  Every call on a device shares it: calls
// that overlap on the GPU (different streams, different host threads) race on it; calls on one stream are ordered.
// Nothing in it has to start at zero -- the GEMM writes every element [nsplit][M][d] that add_ln then reads -- so it is
// not cleared; it keeps the last call's partials.
#include <cuda.h>
#include <mutex>
#include "common.cuh"
#include "launch.h"

int itb_gemm_simt(int dtype, int epi, const void* ctl, const void* A, const void* W, const float* bias, void* out,
                  int M, int N, int K, int ldo, cudaStream_t s);
int itb_gemm_tc(int epi, const CUtensorMap* tmW, const CUtensorMap* tmA, const void* ctl, const float* bias, void* out,
                int M, int N, int K, int ldo, int act_f16, int mt, int nsplit, cudaStream_t s);
int itb_make_tmap_16b(CUtensorMap* map, const void* base, int rows, int K, int box_rows, int is_f16);
void itb_tc_plan(int M, int N, int K, int max_split, int* mt_out, int* split_out);

// [ITB_Y_PARTIALS][M][d] floats, M * d <= ITB_STATIC_SCRATCH_FLOATS / ITB_Y_PARTIALS (8 MiB in all).
static __device__ __align__(256) float g_static_y[ITB_STATIC_SCRATCH_FLOATS];

// In host code `g_static_y` is only a handle for the runtime's symbol table, not a device address: the address is looked
// up once per device, with that device current, and cached by ordinal.
static float* static_y_of_current_device() {
    static float* ptr[ITB_MAX_DEVICES] = {};
    static std::mutex mu;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= ITB_MAX_DEVICES) {
        itb_set_error("linear_add_ln_static: no current device (or ordinal >= %d)", ITB_MAX_DEVICES);
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (ptr[dev] == nullptr) {
        float* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_static_y);
        if (e != cudaSuccess || p == nullptr) {
            itb_set_error("linear_add_ln_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            return nullptr;
        }
        ptr[dev] = p;
    }
    return ptr[dev];
}

extern "C" {

int itb_static_scratch_floats(void) { return ITB_STATIC_SCRATCH_FLOATS; }

int itb_linear_add_ln_static(int dtype, int impl, const void* A, const void* W, const float* bias, const float* g,
                             const float* b, float eps, float* x_f32, void* x_act, int M, int d, int K, void* stream) {
    ITB_REQUIRE(dtype == ITB_F32 || dtype == ITB_BF16 || dtype == ITB_F16, "linear_add_ln_static: unknown dtype");
    ITB_REQUIRE(!(impl == ITB_GEMM_TCGEN05 && dtype == ITB_F32), "linear_add_ln_static: the tcgen05 path needs 16-bit operands");
    ITB_REQUIRE(M > 0 && d > 0 && K > 0, "linear_add_ln_static: bad sizes");
    if ((long long)M * d > ITB_STATIC_SCRATCH_FLOATS / ITB_Y_PARTIALS) {
        itb_set_error("linear_add_ln_static: M * d = %lld exceeds the static scratch (%d floats = %d split-K partials of at most "
                      "%d; use itb_linear + itb_add_ln with your own buffer)", (long long)M * d, ITB_STATIC_SCRATCH_FLOATS,
                      ITB_Y_PARTIALS, ITB_STATIC_SCRATCH_FLOATS / ITB_Y_PARTIALS);
        return -2;
    }
    float* y = static_y_of_current_device();
    if (y == nullptr) return -1;
    cudaStream_t s = (cudaStream_t)stream;
    int nsplit = 1, r;
    if (impl != ITB_GEMM_SIMT && dtype != ITB_F32) {
        int mt = 32;
        CUtensorMap tmW, tmA;
        itb_tc_plan(M, d, K, ITB_Y_PARTIALS, &mt, &nsplit);
        r = itb_make_tmap_16b(&tmW, W, d, K, 128, dtype == ITB_F16 ? 1 : 0);
        if (r == 0) r = itb_make_tmap_16b(&tmA, A, M, K, mt, dtype == ITB_F16 ? 1 : 0);
        if (r == 0) r = itb_gemm_tc(ITB_EPI_BIAS_F32, &tmW, &tmA, nullptr, bias, y, M, d, K, d, dtype == ITB_F16 ? 1 : 0, mt, nsplit, s);
    } else {
        r = itb_gemm_simt(dtype, ITB_EPI_BIAS_F32, nullptr, A, W, bias, y, M, d, K, d, s);
    }
    if (r != 0) return r;
    return itb_add_ln(dtype, nullptr, y, nsplit, g, b, eps, x_f32, x_act, M, d, stream);
}

}  // extern "C"
