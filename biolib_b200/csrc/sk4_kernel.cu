// sk4_kernel.cu -- host side of the fourth-generation fused minimizer / super-k-mer kernel (device code and the
// description of the scheme: sk4_device.cuh): which (k, m) it serves, the launch geometry, the launch.
#include "sk4_device.cuh"

namespace blg {

namespace {

struct Variant4 { int B, Q, LW, MW; };

// w = Q B + 1: B = 12 or 10 with Q = 1 (k=31 m=19, k=21 m=11 and their neighbours), B = 16 with Q = 2 (k=63 m=31), any m: MW <=> the m-mer
// needs a second 32-bit word (m > 16); the block-local streams hold B + m - 1 bases: 2 (B - 1 + m) <= 32 LW.
bool pick_variant4(int k, int m, Variant4* v)
{
    const int d = k - m;
    if (m < 1 || m > 32 || k > 64) return false;
    if (d == 12 || d == 10) { *v = {d, 1, 2 * (d - 1 + m) <= 64 ? 2 : 3, m > 16 ? 1 : 0}; return true; }
    if (d == 32) { *v = {16, 2, 2 * (15 + m) <= 64 ? 2 : 3, m > 16 ? 1 : 0}; return true; }       // k=63 m=31 (config 4): w = 2 * 16 + 1
    return false;
}

constexpr int NB4 = 2;       // blocks per lane (sk4_device.cuh numbers record slots for two)

template <int B, int NB, int Q, int LW, bool CANON, bool MW>
cudaError_t setup4_uncached(size_t smem, int* occ)
{
    cudaError_t e = cudaFuncSetAttribute(sk4_kernel<B, NB, Q, LW, CANON, MW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, sk4_kernel<B, NB, Q, LW, CANON, MW>, SK4_THREADS, smem);
}

// the shared-memory opt-in and the occupancy query are paid once per device and instance, not per call
template <int B, int NB, int Q, int LW, bool CANON, bool MW>
cudaError_t setup4(size_t smem, int* occ)
{
    static int cached_occ[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < 64 && cached_occ[dev] > 0) { *occ = cached_occ[dev]; return cudaSuccess; }
    cudaError_t e = setup4_uncached<B, NB, Q, LW, CANON, MW>(smem, occ);
    if (e == cudaSuccess && dev >= 0 && dev < 64) cached_occ[dev] = *occ;
    return e;
}

template <int B, int NB, int Q, int LW, bool MW>
cudaError_t setup4c(bool canon, size_t smem, int* occ)
{
    return canon ? setup4<B, NB, Q, LW, true, MW>(smem, occ) : setup4<B, NB, Q, LW, false, MW>(smem, occ);
}

template <int B, int NB, int Q, int LW, bool MW>
void launch4c(bool canon, const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream)
{
    if (canon) sk4_kernel<B, NB, Q, LW, true, MW><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    else sk4_kernel<B, NB, Q, LW, false, MW><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
}

template <int B, int Q>
size_t smem4() { return (size_t)Geo4<B, NB4, Q>::WARP_BYTES * WARPS4; }

// (LW, MW) in {(2, false), (2, true), (3, true)}
template <int B, int Q>
cudaError_t setup4v(const Variant4& v, bool canon, size_t smem, int* occ)
{
    if (!v.MW) return setup4c<B, NB4, Q, 2, false>(canon, smem, occ);
    return v.LW == 2 ? setup4c<B, NB4, Q, 2, true>(canon, smem, occ) : setup4c<B, NB4, Q, 3, true>(canon, smem, occ);
}
template <int B, int Q>
void launch4v(const Variant4& v, bool canon, const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream)
{
    if (!v.MW) launch4c<B, NB4, Q, 2, false>(canon, args, plan, stream);
    else if (v.LW == 2) launch4c<B, NB4, Q, 2, true>(canon, args, plan, stream);
    else launch4c<B, NB4, Q, 3, true>(canon, args, plan, stream);
}

}  // namespace

bool sk4_supported(int k, int m)
{
    Variant4 v;
    return pick_variant4(k, m, &v);
}

cudaError_t sk4_plan(SkArgs& args, int sm_count, SkLaunchPlan* plan)
{
    Variant4 v;
    if (!pick_variant4(args.k, args.m, &v)) return cudaErrorInvalidValue;
    int R = 0;
    size_t smem = 0;
    int occ = 0;
    cudaError_t e = cudaSuccess;
    const bool canon = args.canonical != 0;
    if (v.B == 12) { R = Geo4<12, NB4, 1>::R; smem = smem4<12, 1>(); e = setup4v<12, 1>(v, canon, smem, &occ); }
    else if (v.B == 10) { R = Geo4<10, NB4, 1>::R; smem = smem4<10, 1>(); e = setup4v<10, 1>(v, canon, smem, &occ); }
    else { R = Geo4<16, NB4, 2>::R; smem = smem4<16, 2>(); e = setup4v<16, 2>(v, canon, smem, &occ); }
    if (e != cudaSuccess) return e;
    args.R = R;
    args.org = args.win_lo;
    args.keymask = ~63u;
    const long long nwin = args.win_hi > args.org ? args.win_hi - args.org : 0;
    args.num_tiles = args.win_hi > args.win_lo ? (nwin + R - 1) / R : 0;
    args.status_words = args.num_tiles + 2 * ((args.num_tiles + 31) / 32) + 8;
    plan->threads = SK4_THREADS;
    plan->smem_bytes = smem;
    if (occ < 1) occ = 1;
    long long grid = (long long)sm_count * occ;
    const long long units = (args.num_tiles + WARPS4 - 1) / WARPS4;
    if (grid > units) grid = units;
    plan->grid = (int)(grid < 1 ? 1 : grid);
    return cudaSuccess;
}

unsigned int sk4_ticket_draws(const SkArgs& args, const SkLaunchPlan& plan)
{
    return (unsigned int)((unsigned long long)args.num_tiles + (unsigned long long)plan.grid * WARPS4);
}

cudaError_t sk4_launch(const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream)
{
    Variant4 v;
    if (!pick_variant4(args.k, args.m, &v)) return cudaErrorInvalidValue;
    const bool canon = args.canonical != 0;
    if (v.B == 12) launch4v<12, 1>(v, canon, args, plan, stream);
    else if (v.B == 10) launch4v<10, 1>(v, canon, args, plan, stream);
    else launch4v<16, 2>(v, canon, args, plan, stream);
    return cudaGetLastError();
}

}  // namespace blg
