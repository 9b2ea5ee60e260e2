// Host-side helpers shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include "../../include/itb.h"

void itb_set_error(const char* fmt, ...);
int itb_check_launch(const char* what);

#define ITB_REQUIRE(cond, msg)            \
    do {                                  \
        if (!(cond)) {                    \
            itb_set_error("%s", msg);     \
            return -2;                    \
        }                                 \
    } while (0)

// Launches `kern` with programmatic dependent launch enabled (see pdl_wait / pdl_trigger in common.cuh) unless the
// environment sets ITB_PDL=0.  Works in eager streams and under stream capture (the edge becomes programmatic).
bool itb_pdl_enabled();

// Per-device caches.  The reference's own multi-GPU mode is ONE process with a thread per GPU (src/parallel.py:104-148),
// so nothing that CUDA keeps per device (function attributes, SM count) may be cached per process: index by ordinal.
#define ITB_MAX_DEVICES 64
static inline int itb_cur_device() {
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= ITB_MAX_DEVICES) d = 0;
    return d;
}
// SM count of the CURRENT device.
static inline int itb_num_sms() {
    static int n[ITB_MAX_DEVICES] = {};
    const int dev = itb_cur_device();
    if (n[dev] == 0) {
        int v = 0;
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
        n[dev] = v > 0 ? v : 148;
    }
    return n[dev];
}
// Sets MaxDynamicSharedMemorySize of `kern` once per DEVICE (the attribute is per device and per function).
// `done` is a per-call-site `static bool[ITB_MAX_DEVICES]`.  Racing threads on one device both set it: idempotent.
template <typename K>
static inline cudaError_t itb_smem_attr_once(bool* done, K kern, int bytes) {
    const int dev = itb_cur_device();
    if (done[dev]) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) done[dev] = true;
    return e;
}

template <typename... KArgs, typename... Args>
static inline cudaError_t itb_launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = itb_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
