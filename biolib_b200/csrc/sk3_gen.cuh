// sk3_gen.cuh -- body of sk3_gen{16,12,8,6}.cu: instantiates the general form (Q == 0 in sk3_device.cuh) for one block
// size. MW is fixed to true there: both mask words are applied at run time, so one instance serves every m <= 32.
#pragma once
#include "sk3_device.cuh"
#include "sk3_gen.h"

#define BLG_SK3_GEN_DEFINE(B)                                                                                              \
    namespace blg {                                                                                                        \
    namespace {                                                                                                            \
    template <int LW, bool CANON>                                                                                          \
    cudaError_t gen_setup(size_t smem, int* occ)                                                                           \
    {                                                                                                                      \
        cudaError_t e = cudaFuncSetAttribute(sk3_kernel<B, 0, LW, CANON, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                             (int)smem);                                                                   \
        if (e != cudaSuccess) return e;                                                                                    \
        return cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, sk3_kernel<B, 0, LW, CANON, true>, SK3_THREADS, smem);   \
    }                                                                                                                      \
    }                                                                                                                      \
    cudaError_t sk3_gen##B##_setup(int lw, bool canon, size_t* smem, int* occ)                                             \
    {                                                                                                                      \
        *smem = (size_t)Geo<B, 0>::WARP_BYTES * WARPS3;                                                                    \
        if (lw == 2) return canon ? gen_setup<2, true>(*smem, occ) : gen_setup<2, false>(*smem, occ);                      \
        return canon ? gen_setup<3, true>(*smem, occ) : gen_setup<3, false>(*smem, occ);                                   \
    }                                                                                                                      \
    void sk3_gen##B##_launch(int lw, bool canon, const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream)        \
    {                                                                                                                      \
        if (lw == 2) {                                                                                                     \
            if (canon) sk3_kernel<B, 0, 2, true, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);        \
            else sk3_kernel<B, 0, 2, false, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);             \
        } else {                                                                                                           \
            if (canon) sk3_kernel<B, 0, 3, true, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);        \
            else sk3_kernel<B, 0, 3, false, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);             \
        }                                                                                                                  \
    }                                                                                                                      \
    }
