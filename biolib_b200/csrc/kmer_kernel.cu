// kmer_kernel.cu -- batched wrapper::kmer_view for sm_100a: every valid k-mer as (value, position).
//
// Reference semantics (include/kmer_view.hpp:181-202, 213-234): fwd word = bases packed first-base-most-
// significant (:194), rev = reverse complement (:195), value = canonical ? min(fwd, rev) : fwd (:196, the
// strand tie only occurs when both words are equal), a byte outside AaCcGgTtUu resets the run (:198-200).
// For k > 32 the words are 128-bit and the reverse complement is the CORRECT one (the reference shifts a
// 64-bit operand by >= 64 there: SURVEY K7). The reference's iterator quirks (dropped last k-mer, nullopt
// records, ids: K1-K4) are a pure function of the positions emitted here and are rebuilt by the C++ shim.
//
// Shape: tile = 256 threads x 16 consecutive k-mer starts. Each thread extracts its start state from the
// packed 2-bit stream (no warm-up loop), rolls 16 steps, parks the values in shared memory [step][thread]
// (padded, conflict-free both ways); the tile's valid count goes through the same ticket + decoupled
// look-back as the super-k-mer kernel; the output pass is position-parallel, so clean DNA stores fully
// coalesced 8-byte lanes.
#include "common.cuh"
#include "kernels.h"

namespace blg {

namespace {
constexpr int KT = 256;       // threads
constexpr int KP = 16;        // k-mer starts per thread
constexpr int KTILE = KT * KP;
constexpr int KPAD = KT + 1;  // row pitch of the [step][thread] staging arrays

__host__ __device__ inline size_t kmer_smem_bytes(int nwords, bool wide)
{
    size_t off = (size_t)(nwords + 2 + 1) / 2 * 2 * sizeof(uint2);  // cf, 16-byte multiple
    off += (size_t)KP * KPAD * 8 * (wide ? 2 : 1);                  // values
    off += (size_t)KT * 4 * 2;                                      // VM, OFF
    off += 32 * 4 + 8 + 8;                                          // WSUM, SBASE, STILE
    return off;
}
}  // namespace

template <bool WIDE>
__global__ void __launch_bounds__(KT) kmer_kernel(const KmerArgs a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int k = a.k, km1 = k - 1;
    size_t so = 0;
    uint2* cf = reinterpret_cast<uint2*>(smem_raw);
    so += (size_t)(a.nwords + 2 + 1) / 2 * 2 * sizeof(uint2);
    u64* VLO = reinterpret_cast<u64*>(smem_raw + so);
    so += (size_t)KP * KPAD * 8;
    u64* VHI = VLO;
    if (WIDE) { VHI = reinterpret_cast<u64*>(smem_raw + so); so += (size_t)KP * KPAD * 8; }
    u32* VM = reinterpret_cast<u32*>(smem_raw + so);
    so += KT * 4;
    u32* OFF = reinterpret_cast<u32*>(smem_raw + so);
    so += KT * 4;
    u32* WSUM = reinterpret_cast<u32*>(smem_raw + so);
    so += 32 * 4;
    u64* SBASE = reinterpret_cast<u64*>(smem_raw + so);
    so += 8;
    int* STILE = reinterpret_cast<int*>(smem_raw + so);

    const bool canon = a.canonical != 0;
    typedef unsigned __int128 u128;

    for (;;) {
        if (tid == 0) STILE[0] = (int)atomicAdd(a.tile_counter, 1u);
        __syncthreads();
        const long long tile = (unsigned)STILE[0];
        if (tile >= a.num_tiles) break;
        const long long tile_lo = a.win_lo + tile * (long long)KTILE;
        const long long A0 = tile_lo & ~15LL;
        const int off0 = (int)(tile_lo - A0);

        stage_tile(cf, a.nwords, a.bases, a.buf_lo, a.buf_hi, A0, a.vec_ok, tid, KT);
        __syncthreads();
        if (a.seq_offsets != nullptr) {
            if (warp == 0) mark_sequence_starts(cf, a.nwords, a.seq_offsets, a.n_seq, A0, lane);
            __syncthreads();
        }

        const int lp = off0 + tid * KP;
        u32 vm = 0;
        {
            // start state from the k-1 bases [lp, lp + k - 1)
            int cnt;
            u64 f0 = stream64_flags(cf, lp), f1 = 0;
            u64 c0 = stream64_codes(cf, lp), c1 = 0;
            if (WIDE) { f1 = stream64_flags(cf, lp + 32); c1 = stream64_codes(cf, lp + 32); }
            if (!WIDE) {
                const u64 mk = km1 ? (~0ULL >> (64 - 2 * km1)) : 0ULL;
                f0 &= mk; c0 &= mk;
            } else {
                const u64 mk = km1 > 32 ? (~0ULL >> (128 - 2 * km1)) : 0ULL;  // km1 >= 32 here
                f1 &= mk; c1 &= mk;
            }
            if ((f0 | f1) == 0) cnt = km1;
            else {
                const int hb = f1 ? 64 + (63 - __clzll((long long)f1)) : (63 - __clzll((long long)f0));
                const int q = hb >> 1;
                const u64 fl = q >= 32 ? (f1 >> (2 * (q - 32))) : (f0 >> (2 * q));
                cnt = km1 - q - (((fl & 3ULL) == 2ULL) ? 0 : 1);
            }
            if (!WIDE) {
                const u64 mk1 = km1 ? (~0ULL >> (64 - 2 * km1)) : 0ULL;
                const u64 kmask = ~0ULL >> (64 - 2 * k);
                u64 rc = ((~c0) & mk1) << 2;
                u64 fwd = km1 ? (pairrev64(c0) >> (64 - 2 * km1)) : 0ULL;
                const int rshift = 2 * km1;
#pragma unroll
                for (int i = 0; i < KP; ++i) {
                    const int q = lp + km1 + i;
                    const uint2 word = cf[q >> 4];
                    const int sh = (q & 15) * 2;
                    const u32 c = (word.x >> sh) & 3u, f = (word.y >> sh) & 3u;
                    fwd = ((fwd << 2) | c) & kmask;
                    rc = (rc >> 2) | ((u64)(3u ^ c) << rshift);
                    cnt = f == 0 ? cnt + 1 : (f == 2 ? 1 : 0);
                    vm |= (u32)(cnt >= k) << i;
                    VLO[i * KPAD + tid] = (canon && rc < fwd) ? rc : fwd;
                }
            } else {
                const u128 chunk = ((u128)c1 << 64) | c0;
                const u128 mk1 = (~(u128)0) >> (128 - 2 * km1);
                const u128 kmask = (~(u128)0) >> (128 - 2 * k);
                u128 rc = ((~chunk) & mk1) << 2;
                u128 fwd = ((((u128)pairrev64(c0)) << 64) | pairrev64(c1)) >> (128 - 2 * km1);
                const int rshift = 2 * km1;
#pragma unroll
                for (int i = 0; i < KP; ++i) {
                    const int q = lp + km1 + i;
                    const uint2 word = cf[q >> 4];
                    const int sh = (q & 15) * 2;
                    const u32 c = (word.x >> sh) & 3u, f = (word.y >> sh) & 3u;
                    fwd = ((fwd << 2) | c) & kmask;
                    rc = (rc >> 2) | ((u128)(3u ^ c) << rshift);
                    cnt = f == 0 ? cnt + 1 : (f == 2 ? 1 : 0);
                    vm |= (u32)(cnt >= k) << i;
                    const u128 v = (canon && rc < fwd) ? rc : fwd;
                    VLO[i * KPAD + tid] = (u64)v;
                    VHI[i * KPAD + tid] = (u64)(v >> 64);
                }
            }
        }
        // windows past win_hi belong to nobody
        {
            long long nwin = a.win_hi - (tile_lo + (long long)tid * KP);
            nwin = nwin < 0 ? 0 : (nwin > KP ? KP : nwin);
            vm &= (1u << nwin) - 1u;
        }
        u32 cntv = __popc(vm), incl = cntv;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const u32 v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) WSUM[warp] = incl;
        __syncthreads();
        u32 warp_off = 0, tile_total = 0;
#pragma unroll
        for (int q = 0; q < KT / 32; ++q) {
            const u32 v = WSUM[q];
            if (q < warp) warp_off += v;
            tile_total += v;
        }
        VM[tid] = vm;
        OFF[tid] = warp_off + incl - cntv;
        if (warp == 0) {
            const u64 excl = lookback_exclusive(a.tile_status, tile, tile_total, lane);
            if (lane == 0) {
                SBASE[0] = excl;
                if (tile == a.num_tiles - 1) *a.total = excl + tile_total;
            }
        }
        __syncthreads();

        const u64 base = SBASE[0];
#pragma unroll 4
        for (int j = 0; j < KP; ++j) {
            const int p = tid + j * KT;          // tile-local k-mer start
            const int t = p >> 4, i = p & 15;    // owner thread, step
            const u32 m_ = VM[t];
            if ((m_ >> i) & 1u) {
                const u64 g = base + OFF[t] + __popc(m_ & ((1u << i) - 1u));
                if (g < a.capacity) {
                    a.value_lo[g] = VLO[i * KPAD + t];
                    if (WIDE) a.value_hi[g] = VHI[i * KPAD + t];
                    if (a.position != nullptr) a.position[g] = (u64)(tile_lo + p);
                }
            }
        }
        __syncthreads();
    }
}

cudaError_t kmer_plan(KmerArgs& args, int sm_count, KmerLaunchPlan* plan)
{
    const bool wide = args.k > 32;
    args.nwords = (KTILE + args.k + 30) / 16 + 1;
    const long long nwin = args.win_hi > args.win_lo ? args.win_hi - args.win_lo : 0;
    args.num_tiles = (nwin + KTILE - 1) / KTILE;
    plan->threads = KT;
    plan->smem_bytes = kmer_smem_bytes(args.nwords, wide);
    cudaError_t e;
    int occ = 0;
    if (wide) {
        e = cudaFuncSetAttribute(kmer_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan->smem_bytes);
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kmer_kernel<true>, KT, plan->smem_bytes);
    } else {
        e = cudaFuncSetAttribute(kmer_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan->smem_bytes);
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kmer_kernel<false>, KT, plan->smem_bytes);
    }
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
    long long grid = (long long)sm_count * occ;
    if (grid > args.num_tiles) grid = args.num_tiles;
    plan->grid = (int)(grid < 1 ? 1 : grid);
    return cudaSuccess;
}

cudaError_t kmer_launch(const KmerArgs& args, const KmerLaunchPlan& plan, cudaStream_t stream)
{
    if (args.k > 32) kmer_kernel<true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    else kmer_kernel<false><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    return cudaGetLastError();
}

}  // namespace blg
