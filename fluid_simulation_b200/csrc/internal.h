// internal.h -- entry points shared between the translation units of the library (api.cpp,
// kernels.cu) that are not part of the C ABI in include/fluidb200.h.
#pragma once
#include "../../include/fluidb200.h"

namespace fbapi {

enum {
    POISSON_NO_SPLIT = 1,       // one kernel does every task: no helper stream, no events
    POISSON_VALIDATE_ONLY = 2,  // check the arguments like a launch would, launch nothing
};
// fluidb200_poisson_sweeps with options (same arguments, same validation, same error reporting).
int poisson_sweeps(const fluidb200_slab* g, const void* p_in, void* p_out, const void* rhs,
                   const fluidb200_edges* p_edges, const fluidb200_mask* mask, int sweeps, int out_lo,
                   int out_hi, void* stream, int flags);

}  // namespace fbapi
