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
