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
