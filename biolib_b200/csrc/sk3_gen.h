// sk3_gen.h -- the general form of the sk3 kernel (any k - m = q B + r), one block size B per translation unit so that
// the sixteen instantiations (B x stream words x canonical) compile in parallel. Called by sk3_kernel.cu only.
#pragma once
#include <cuda_runtime.h>
#include "kernels.h"

namespace blg {

// lw = 32-bit words of the lane-local streams (2 or 3; 2 (m + B - 1) <= 32 lw). setup: opt-in shared memory + occupancy.
#define BLG_SK3_GEN_DECL(B)                                                                                  \
    cudaError_t sk3_gen##B##_setup(int lw, bool canon, size_t* smem, int* occ);                              \
    void sk3_gen##B##_launch(int lw, bool canon, const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream);
BLG_SK3_GEN_DECL(16)
BLG_SK3_GEN_DECL(12)
BLG_SK3_GEN_DECL(8)
BLG_SK3_GEN_DECL(6)
#undef BLG_SK3_GEN_DECL

}  // namespace blg
