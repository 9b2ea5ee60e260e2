// Diagnostic build only (tools/timeline.py): every translation unit of libitb.so in ONE, compiled with -DITB_TIMELINE,
// so that the kernels share the stamp buffer of common.cuh.  Build: csrc/build_timeline.sh -> csrc/build/libitb_timeline.so,
// loaded through ITB_LIB.  Not part of the product.
#define ITB_TL_FILE 1
#include "misc.cu"
#undef ITB_TL_FILE
#define ITB_TL_FILE 2
#include "attention.cu"
#undef ITB_TL_FILE
#define ITB_TL_FILE 3
#include "attention_mma.cu"
#undef ITB_TL_FILE
#define ITB_TL_FILE 4
#include "attention_bulk.cu"
#undef ITB_TL_FILE
#define ITB_TL_FILE 5
#include "beam.cu"
#undef ITB_TL_FILE
#define ITB_TL_FILE 6
#include "gemm_simt.cu"
#undef ITB_TL_FILE
#define ITB_TL_FILE 7
#include "gemm_tc.cu"
#undef ITB_TL_FILE
#define ITB_TL_FILE 8
#include "api.cu"
#include "no_alloc.cu"

extern "C" {
// copies the stamps to host: out[0] = count, then (tag, time ns) pairs; returns the count (clamped to cap)
int itb_timeline_read(unsigned long long* out, int cap_pairs) {
    unsigned long long n = 0;
    if (cudaMemcpyFromSymbol(&n, g_itb_tl, sizeof(n)) != cudaSuccess) return -1;
    if (n > ITB_TL_CAP) n = ITB_TL_CAP;
    if ((int)n > cap_pairs) n = cap_pairs;
    out[0] = n;
    if (n > 0 && cudaMemcpyFromSymbol(out + 1, g_itb_tl, 2 * n * sizeof(unsigned long long), sizeof(unsigned long long)) != cudaSuccess) return -1;
    return (int)n;
}
int itb_timeline_reset(void) {
    const unsigned long long z = 0;
    return cudaMemcpyToSymbol(g_itb_tl, &z, sizeof(z)) == cudaSuccess ? 0 : -1;
}
}
