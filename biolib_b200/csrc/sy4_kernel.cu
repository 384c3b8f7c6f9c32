// sy4_kernel.cu -- host side of the syncmer filter in the sk4 frame (device code: sy4_device.cuh): which (k, m) it serves,
// the launch geometry, the launch. Everything else -- k - m other than 12 or 10 -- stays with kmer2_kernel<.., SYNC>.
#include "sy4_device.cuh"

namespace blg {

namespace {

struct VariantS { int B, LW, MW; };

bool pick_variant_s(int k, int m, VariantS* v)
{
    const int d = k - m;
    if (m < 1 || k > 32 || (d != 12 && d != 10)) return false;
    *v = {d, 2 * (d - 1 + m) <= 64 ? 2 : 3, m > 16 ? 1 : 0};
    return true;
}

constexpr int NBS = 2;

template <int B, int LW, bool CANON, bool MW>
cudaError_t setup_s(size_t smem, int* occ)
{
    static int cached_occ[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < 64 && cached_occ[dev] > 0) { *occ = cached_occ[dev]; return cudaSuccess; }
    cudaError_t e = cudaFuncSetAttribute(sy4_kernel<B, NBS, LW, CANON, MW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, sy4_kernel<B, NBS, LW, CANON, MW>, SK4_THREADS, smem);
    if (e == cudaSuccess && dev >= 0 && dev < 64) cached_occ[dev] = *occ;
    return e;
}

template <int B, bool CANON>
size_t smem_s() { return (size_t)Geo4<B, NBS, 1, CANON ? 2 : 1>::WARP_BYTES * WARPS4; }

template <int B, bool CANON>
cudaError_t setup_v(const VariantS& v, size_t smem, int* occ)
{
    if (!v.MW) return setup_s<B, 2, CANON, false>(smem, occ);
    return v.LW == 2 ? setup_s<B, 2, CANON, true>(smem, occ) : setup_s<B, 3, CANON, true>(smem, occ);
}

template <int B, bool CANON>
void launch_v(const VariantS& v, const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream)
{
    if (!v.MW) sy4_kernel<B, NBS, 2, CANON, false><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    else if (v.LW == 2) sy4_kernel<B, NBS, 2, CANON, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    else sy4_kernel<B, NBS, 3, CANON, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
}

}  // namespace

bool sy4_supported(int k, int m)
{
    VariantS v;
    return pick_variant_s(k, m, &v);
}

cudaError_t sy4_plan(SkArgs& args, int sm_count, SkLaunchPlan* plan)
{
    VariantS v;
    if (!pick_variant_s(args.k, args.m, &v)) return cudaErrorInvalidValue;
    const bool canon = args.canonical != 0;
    int R = 0, occ = 0;
    size_t smem = 0;
    cudaError_t e;
    if (v.B == 12) {
        R = Geo4<12, NBS, 1>::R;
        if (canon) { smem = smem_s<12, true>(); e = setup_v<12, true>(v, smem, &occ); } else { smem = smem_s<12, false>(); e = setup_v<12, false>(v, smem, &occ); }
    } else {
        R = Geo4<10, NBS, 1>::R;
        if (canon) { smem = smem_s<10, true>(); e = setup_v<10, true>(v, smem, &occ); } else { smem = smem_s<10, false>(); e = setup_v<10, false>(v, smem, &occ); }
    }
    if (e != cudaSuccess) return e;
    args.R = R;
    args.org = args.win_lo;
    args.keymask = ~63u;
    const long long nwin = args.win_hi > args.org ? args.win_hi - args.org : 0;
    args.num_tiles = (nwin + R - 1) / R;
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

cudaError_t sy4_launch(const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream)
{
    VariantS v;
    if (!pick_variant_s(args.k, args.m, &v)) return cudaErrorInvalidValue;
    const bool canon = args.canonical != 0;
    if (v.B == 12) { if (canon) launch_v<12, true>(v, args, plan, stream); else launch_v<12, false>(v, args, plan, stream); }
    else { if (canon) launch_v<10, true>(v, args, plan, stream); else launch_v<10, false>(v, args, plan, stream); }
    return cudaGetLastError();
}

}  // namespace blg
