// sk_kernel.cu -- fused minimizer / super-k-mer extraction for sm_100a.
//
// What it computes (decisions D1-D8; reference intent: include/minimizer_view.hpp:184-266,
// include/super_kmer_view.hpp:121-135): for every valid k-mer start s the argmin over its w = k-m+1 m-mers
// of (hash64(value, seed), position) -- smallest hash, leftmost on ties -- and one record per maximal run of
// consecutive valid k-mers that share the same argmin POSITION: {value, mm_start, size, mm_pos}.
//
// How (not a translation of the iterators -- nothing is rolled across threads):
//   tile    = blockDim.x consecutive "w-blocks" of w positions each; thread t owns block t-1 of the tile
//             (block -1 and the two blocks after the last real one are ghosts that only feed neighbours,
//             so every record that starts in a real block is finished inside the tile: the k+w-2 halo).
//   stage   ASCII -> 2-bit codes + break flags in shared memory, 16-byte coalesced loads, SWAR table.
//   phase A each thread rolls fwd / reverse-complement words over its w m-mers (start state extracted
//           directly from the packed stream, no warm-up loop), hashes each m-mer exactly once
//           (MurmurHash3 x64_128, 8-byte key) and keeps the running prefix minimum of its block.
//   phase B van Herk / Gil-Werman: running suffix minimum of the own block combined with the prefix
//           minimum of the right neighbour gives the window argmin of each of the thread's w windows;
//           transitions between consecutive windows become boundary bits (one 64-bit mask per thread).
//   phase C popcount of the start bits, block scan, decoupled look-back across tiles (ordered output).
//   phase D each start bit becomes a 32-bit record descriptor in shared memory (size from bit scans).
//   phase E record-parallel: rebuild the minimizer value from the packed stream and write the SoA
//           outputs with fully coalesced stores.
// All per-(block, i) shared arrays are laid out [i][thread] so that lanes touch consecutive words.
#include "common.cuh"
#include "kernels.h"

namespace blg {

namespace {

struct SkSmem {
    uint2* cf;    // [nwords + 2]
    u64* H;       // [w][NB] hashes; reused as u32 REC[] once phase B is done
    u64* PRE;     // [w][NB] prefix-min hash of the block up to i
    u8* PREI;     // [w][NB] its offset | invalid-so-far << 7
    u8* AMIN;     // [w][NB] window argmin offset from block start (0..2w-2) | window-invalid << 7
    u8* LAST;     // [NB]    AMIN of the block's last window
    u64* BM;      // [NB]    boundary-or-invalid mask of the block's windows
    u32* WSUM;    // [32]
    u64* SBASE;   // [1]
    int* STILE;   // [2]
};

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

__host__ __device__ inline size_t sk_carve(SkSmem* s, unsigned char* base, int nwords, int w, int NB)
{
    size_t off = 0;
    if (s) s->cf = reinterpret_cast<uint2*>(base + off);
    off += align_up((size_t)(nwords + 2) * sizeof(uint2), 16);
    if (s) s->H = reinterpret_cast<u64*>(base + off);
    off += (size_t)w * NB * 8;
    if (s) s->PRE = reinterpret_cast<u64*>(base + off);
    off += (size_t)w * NB * 8;
    if (s) s->PREI = base + off;
    off += (size_t)w * NB;
    if (s) s->AMIN = base + off;
    off += (size_t)w * NB;
    if (s) s->LAST = base + off;
    off += align_up((size_t)NB, 8);
    off = align_up(off, 8);
    if (s) s->BM = reinterpret_cast<u64*>(base + off);
    off += (size_t)NB * 8;
    if (s) s->WSUM = reinterpret_cast<u32*>(base + off);
    off += 32 * 4;
    if (s) s->SBASE = reinterpret_cast<u64*>(base + off);
    off += 8;
    if (s) s->STILE = reinterpret_cast<int*>(base + off);
    off += 8;
    return off;
}

}  // namespace

__global__ void __launch_bounds__(256) sk_kernel(const SkArgs a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int NB = blockDim.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int w = a.w, m = a.m;
    SkSmem S;
    sk_carve(&S, smem_raw, a.nwords, w, NB);
    u32* REC = reinterpret_cast<u32*>(S.H);

    const int mm1 = m - 1;
    const u64 mask1 = mm1 ? (~0ULL >> (64 - 2 * mm1)) : 0ULL;   // m-1 bases
    const u64 mmask = ~0ULL >> (64 - 2 * m);                    // m bases (m <= 32)
    const int rshift = 2 * mm1;
    const bool canon = a.canonical != 0;
    const u64 s8 = a.s8;
    const long long tile_span = (long long)a.rb * w;

    for (;;) {
        // ---- ticket -------------------------------------------------------------------------
        if (tid == 0) {
            const unsigned t = atomicAdd(a.tile_counter, 1u);
            S.STILE[0] = (int)t;
        }
        __syncthreads();
        const long long tile = (unsigned)S.STILE[0];
        if (tile >= a.num_tiles) break;

        const long long tile_lo = a.win_lo + tile * tile_span;
        const long long A0 = (tile_lo - w) & ~15LL;
        const int off0 = (int)(tile_lo - A0);  // stream index of window 0 of the tile (>= w)

        // ---- stage --------------------------------------------------------------------------
        stage_tile(S.cf, a.nwords, a.bases, a.buf_lo, a.buf_hi, A0, a.vec_ok, tid, NB);
        __syncthreads();
        if (a.seq_offsets != nullptr) {
            if (warp == 0) mark_sequence_starts(S.cf, a.nwords, a.seq_offsets, a.n_seq, A0, lane);
            __syncthreads();
        }

        // ---- phase A: roll + hash + prefix minima -----------------------------------------------
        const int lp = off0 + (tid - 1) * w;  // stream index of this thread's first m-mer
        u64 minv = 0;                         // bit i: m-mer i of the block is invalid
        u64 smask = 0;                        // m == 1 only: bit i: a sequence starts at m-mer i (no m-mer straddles it)
        {
            u64 chunk = stream64_codes(S.cf, lp) & mask1;
            const u64 fchunk = stream64_flags(S.cf, lp) & mask1;
            u64 rc = ((~chunk) & mask1) << 2;
            u64 fwd = mm1 ? (pairrev64(chunk) >> (64 - 2 * mm1)) : 0ULL;
            int cnt;  // valid bases ending at the current one, inside one sequence (saturates harmlessly)
            if (fchunk == 0) cnt = mm1;
            else {
                const int q = (63 - __clzll((long long)fchunk)) >> 1;
                cnt = mm1 - q - ((((fchunk >> (2 * q)) & 3ULL) == 2ULL) ? 0 : 1);
            }
            u64 ph = 0;
            u32 pi = 0, pinv = 0;
            for (int i = 0; i < w; ++i) {
                const int q = lp + mm1 + i;
                const uint2 word = S.cf[q >> 4];
                const int sh = (q & 15) * 2;
                const u32 c = (word.x >> sh) & 3u, f = (word.y >> sh) & 3u;
                fwd = ((fwd << 2) | c) & mmask;
                rc = (rc >> 2) | ((u64)(3u ^ c) << rshift);
                cnt = f == 0 ? cnt + 1 : (f == 2 ? 1 : 0);
                const u32 inv = cnt < m;
                const u64 val = (canon && rc < fwd) ? rc : fwd;
                const u64 h = murmur3_key8(val, s8);
                if (i == 0 || h < ph) { ph = h; pi = i; }   // strict: leftmost wins (minimizer_view.hpp:283)
                pinv |= inv;
                minv |= (u64)inv << i;
                if (m == 1) { const u32 st = f >> 1; smask |= (u64)st << i; pinv |= st; }
                S.H[i * NB + tid] = h;
                S.PRE[i * NB + tid] = ph;
                S.PREI[i * NB + tid] = (u8)(pi | (pinv << 7));
            }
        }
        __syncthreads();

        // ---- phase B: suffix minima + neighbour prefix -> window argmin, boundary bits ------------
        u64 bmask = 0, vmask = 0;
        u32 first_am = 0, first_inv = 1;
        if (tid < NB - 1) {
            u64 sh_ = 0;
            u32 si = 0, sinv = 0, nam = 0, ninv = 1;
            for (int i = w - 1; i >= 0; --i) {
                const u64 h = S.H[i * NB + tid];
                if (i == w - 1 || h <= sh_) { sh_ = h; si = i; }   // '<=': the left element wins ties
                sinv |= (u32)(minv >> i) & 1u;
                u32 am = si, winv = sinv | (u32)((smask >> i) >> 1 != 0);   // a sequence start after window start i
                if (i > 0) {
                    const u64 ph = S.PRE[(i - 1) * NB + tid + 1];
                    const u32 pb = S.PREI[(i - 1) * NB + tid + 1];
                    winv |= pb >> 7;
                    if (ph < sh_) am = w + (pb & 63u);             // strict: the suffix (left) side wins ties
                }
                S.AMIN[i * NB + tid] = (u8)(am | (winv << 7));
                if (i == w - 1) S.LAST[tid] = (u8)(am | (winv << 7));
                else bmask |= (u64)(winv | ninv | (am != nam)) << (i + 1);
                vmask |= (u64)(winv ^ 1u) << i;
                nam = am; ninv = winv;
            }
            first_am = nam; first_inv = ninv;
        }
        __syncthreads();

        // ---- phase C: first boundary bit, start mask, ordered offsets ---------------------------
        u64 startmask = 0;
        if (tid >= 1 && tid < NB - 1) {
            const u32 l = S.LAST[tid - 1];
            bmask |= (u64)(first_inv | (l >> 7) | ((l & 127u) != first_am + (u32)w));
        } else {
            bmask |= 1ULL;
        }
        if (tid < NB - 1) S.BM[tid] = (bmask | ~vmask) & lowmask64(w);
        if (tid >= 1 && tid <= a.rb) {
            const long long s_first = tile_lo + (long long)(tid - 1) * w;
            long long nwin = a.win_hi - s_first;
            nwin = nwin < 0 ? 0 : (nwin > w ? w : nwin);
            startmask = bmask & vmask & lowmask64((int)nwin);
        }
        u32 cnt = __popcll(startmask), incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const u32 v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) S.WSUM[warp] = incl;
        __syncthreads();
        u32 warp_off = 0, tile_total = 0;
        for (int q = 0; q < (NB >> 5); ++q) {
            const u32 v = S.WSUM[q];
            if (q < warp) warp_off += v;
            tile_total += v;
        }
        u32 off = warp_off + incl - cnt;
        if (warp == 0) {
            const u64 excl = lookback_exclusive(a.tile_status, tile, tile_total, lane);
            if (lane == 0) {
                S.SBASE[0] = excl;
                if (tile == a.num_tiles - 1) *a.total = excl + tile_total;
            }
        }

        // ---- phase D: record descriptors --------------------------------------------------------
        if (startmask) {
            const u64 own = S.BM[tid], next = S.BM[tid + 1];
            u64 sm = startmask;
            while (sm) {
                const int i = __ffsll((long long)sm) - 1;
                sm &= sm - 1;
                const u64 above = i == 63 ? 0ULL : (own >> (i + 1));
                int size = above ? __ffsll((long long)above) : (w - i) + (next ? __ffsll((long long)next) - 1 : w);
                size = size > w ? w : size;
                const u32 am = S.AMIN[i * NB + tid] & 127u;
                REC[off++] = (u32)((tid - 1) * w + i) | ((am - (u32)i) << 14) | ((u32)size << 20);
            }
        }
        __syncthreads();

        // ---- phase E: record-parallel output ------------------------------------------------------
        {
            const u64 base = S.SBASE[0];
            for (u32 r = tid; r < tile_total; r += NB) {
                const u32 rec = REC[r];
                const u32 sl = rec & 0x3FFFu, mp = (rec >> 14) & 63u, sz = rec >> 20;
                const u32 jl = sl + mp;
                const u64 chunk = stream64_codes(S.cf, off0 + (int)jl) & mmask;
                const u64 rcw = (~chunk) & mmask;
                const u64 fw = pairrev64(chunk) >> (64 - 2 * m);
                const u64 val = (canon && rcw < fw) ? rcw : fw;
                const u64 g = base + r;
                if (g < a.capacity) {
                    a.value[g] = val;
                    a.mm_start[g] = (u64)(tile_lo + jl);
                    if (a.size != nullptr) { a.size[g] = (u8)sz; a.mm_pos[g] = (u8)mp; }
                    if (a.hash != nullptr) a.hash[g] = murmur3_key8(val, s8);
                }
            }
        }
        __syncthreads();  // REC / cf are rewritten by the next tile
    }
}

cudaError_t sk_plan(SkArgs& args, int sm_count, SkLaunchPlan* plan)
{
    const int w = args.w, m = args.m;
    // largest CTA (multiple of 32, 64..256) whose tile fits ~72 KB so that three CTAs share an SM
    int NB = 256;
    while (NB > 64 && sk_carve(nullptr, nullptr, (NB * w + m + 30) / 16 + 1, w, NB) > 72 * 1024) NB -= 32;
    args.rb = NB - 3;
    args.nwords = (NB * w + m + 30) / 16 + 1;
    const long long span = (long long)args.rb * w;
    const long long nwin = args.win_hi > args.win_lo ? args.win_hi - args.win_lo : 0;
    args.num_tiles = (nwin + span - 1) / span;
    plan->threads = NB;
    plan->smem_bytes = sk_carve(nullptr, nullptr, args.nwords, w, NB);
    cudaError_t e = cudaFuncSetAttribute(sk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan->smem_bytes);
    if (e != cudaSuccess) return e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sk_kernel, NB, plan->smem_bytes);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
    long long grid = (long long)sm_count * occ;
    if (grid > args.num_tiles) grid = args.num_tiles;
    plan->grid = (int)(grid < 1 ? 1 : grid);
    return cudaSuccess;
}

cudaError_t sk_launch(const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream)
{
    sk_kernel<<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    return cudaGetLastError();
}

// ---- batched hash::hash64::hash<uint64_t> (include/hash.hpp:55-59) ------------------------------
// Memory-bound (16 B per value against 40 integer instructions): two values per 16-byte access, four accesses in
// flight per thread before the first hash is computed; persistent grid of 148 x 8 CTAs.
__global__ void __launch_bounds__(256) hash64_kernel(const u64* __restrict__ in, u64* __restrict__ out, size_t n, u64 s8)
{
    const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
    const bool vec = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0;
    size_t done = 0;
    if (vec) {
        const ulonglong2* in2 = reinterpret_cast<const ulonglong2*>(in);
        ulonglong2* out2 = reinterpret_cast<ulonglong2*>(out);
        const size_t n2 = n / 2;
        size_t i = tid;
        for (; i + 3 * nth < n2; i += 4 * nth) {
            ulonglong2 v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = __ldg(in2 + i + j * nth);
#pragma unroll
            for (int j = 0; j < 4; ++j) out2[i + j * nth] = make_ulonglong2(murmur3_key8(v[j].x, s8), murmur3_key8(v[j].y, s8));
        }
        for (; i < n2; i += nth) {
            const ulonglong2 v = __ldg(in2 + i);
            out2[i] = make_ulonglong2(murmur3_key8(v.x, s8), murmur3_key8(v.y, s8));
        }
        done = n2 * 2;
    }
    for (size_t i = done + tid; i < n; i += nth) out[i] = murmur3_key8(in[i], s8);
}

cudaError_t hash64_launch(const unsigned long long* in, unsigned long long* out, size_t n, unsigned long long s8,
                          cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    size_t blocks = (n / 2 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    hash64_kernel<<<(unsigned)blocks, 256, 0, stream>>>(in, out, n, s8);
    return cudaGetLastError();
}

}  // namespace blg
