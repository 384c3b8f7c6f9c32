// sk3_kernel.cu -- host side of the third-generation fused minimizer / super-k-mer kernel (device code and the
// description of the scheme: sk3_device.cuh): which form serves a (k, m), the launch geometry, the launch.
//
//   general form  (k - m = q B + r, q and r launch arguments)   every other (k, m) with k - m >= 5, compiled one block
//                                                               size per translation unit in sk3_gen{16,12,8,6}.cu
#include "sk3_device.cuh"
#include "sk3_gen.h"

namespace blg {

// ---- host side: which (k, m) this generation covers, and the launch -------------------------------------------------
namespace {

struct Variant { int B, Q, LW, MW, gq, gr, gl; };   // Q == 0: general form with k - m = gq B + gr, lanes 1 .. gl own starts

// Tuned: k - m = Q * B. General: every window must be suffix-of-own-block + whole blocks + prefix-of-a-block, which
// holds for any k - m >= B and for k - m = B - 1. Both: m-mer + block must fit the lane-local stream,
// 2 (m + B - 1) <= 32 LW; MW <=> m > 16 (tuned forms only).
bool pick_variant(int k, int m, Variant* v)
{
    const int d = k - m;
    if (m < 1 || m > 32 || k > 64) return false;
    // the tuned forms (k - m = 12, 10, 32: configs 2 - 5) were superseded by sk4_kernel.cu: only the general form is left
    const int B = d >= 15 ? 16 : d >= 11 ? 12 : d >= 7 ? 8 : d >= 5 ? 6 : 0;
    if (B == 0) return false;                                                            // w <= 5: sk2
    const int q = d / B, r = d % B;
    const int gl = 31 - 2 * q - (r ? 2 : 0);
    if (gl < 1) return false;
    *v = {B, 0, 2 * (m + B - 1) <= 64 ? 2 : 3, 1, q, r, gl};
    return true;
}

template <int B, int Q, int LW, bool CANON, bool MW>
cudaError_t setup3(size_t smem, int* occ)
{
    cudaError_t e = cudaFuncSetAttribute(sk3_kernel<B, Q, LW, CANON, MW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, sk3_kernel<B, Q, LW, CANON, MW>, SK3_THREADS, smem);
}

template <int B, int Q, int LW, bool MW>
cudaError_t setup3c(bool canon, size_t smem, int* occ)
{
    return canon ? setup3<B, Q, LW, true, MW>(smem, occ) : setup3<B, Q, LW, false, MW>(smem, occ);
}

template <int B, int Q, int LW, bool MW>
void launch3c(bool canon, const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream)
{
    if (canon) sk3_kernel<B, Q, LW, true, MW><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    else sk3_kernel<B, Q, LW, false, MW><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
}

template <int B, int Q>
size_t smem3() { return (size_t)Geo<B, Q>::WARP_BYTES * WARPS3; }

}  // namespace

bool sk3_supported(int k, int m)
{
    Variant v;
    return pick_variant(k, m, &v);
}

cudaError_t sk3_plan(SkArgs& args, int sm_count, SkLaunchPlan* plan)
{
    Variant v;
    if (!pick_variant(args.k, args.m, &v)) return cudaErrorInvalidValue;
    int R = 0;
    size_t smem = 0;
    int occ = 0;
    cudaError_t e = cudaSuccess;
    const bool canon = args.canonical != 0;
    if (v.Q == 0) {
        R = v.gl * v.B;
        args.gq = v.gq; args.gr = v.gr; args.gl = v.gl;
        if (v.B == 16) e = sk3_gen16_setup(v.LW, canon, &smem, &occ);
        else if (v.B == 12) e = sk3_gen12_setup(v.LW, canon, &smem, &occ);
        else if (v.B == 8) e = sk3_gen8_setup(v.LW, canon, &smem, &occ);
        else e = sk3_gen6_setup(v.LW, canon, &smem, &occ);
    }
    else return cudaErrorInvalidValue;
    if (e != cudaSuccess) return e;
    args.R = R;
    args.org = args.win_lo;
    const long long nwin = args.win_hi > args.org ? args.win_hi - args.org : 0;
    args.num_tiles = args.win_hi > args.win_lo ? (nwin + R - 1) / R : 0;
    args.status_words = args.num_tiles + 2 * ((args.num_tiles + 31) / 32) + 8;
    plan->threads = SK3_THREADS;
    plan->smem_bytes = smem;
    if (occ < 1) occ = 1;
    long long grid = (long long)sm_count * occ;
    const long long units = (args.num_tiles + WARPS3 - 1) / WARPS3;
    if (grid > units) grid = units;
    plan->grid = (int)(grid < 1 ? 1 : grid);
    return cudaSuccess;
}

cudaError_t sk3_launch(const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream)
{
    Variant v;
    if (!pick_variant(args.k, args.m, &v)) return cudaErrorInvalidValue;
    const bool canon = args.canonical != 0;
    if (v.Q == 0) {
        if (v.B == 16) sk3_gen16_launch(v.LW, canon, args, plan, stream);
        else if (v.B == 12) sk3_gen12_launch(v.LW, canon, args, plan, stream);
        else if (v.B == 8) sk3_gen8_launch(v.LW, canon, args, plan, stream);
        else sk3_gen6_launch(v.LW, canon, args, plan, stream);
    }
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

}  // namespace blg
