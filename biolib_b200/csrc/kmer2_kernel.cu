// kmer2_kernel.cu -- batched wrapper::kmer_view for sm_100a, second generation: every valid k-mer as (value, position).
//
// Same results as kmer_kernel.cu (reference semantics include/kmer_view.hpp:181-202, 213-234). What
// changed is the instruction count: the profile of the first kernel (profiles/r01_s29_kmer_kernel_details.txt) shows 94
// executed instructions per k-mer with the ALU pipe 69 % busy and DRAM at 44 % -- the per-base rolling update of two
// 64-bit words, the [step][thread] transposition through shared memory and the per-element offset lookup cost more than
// the memory traffic. Here
//
//   * nothing rolls: the tile's bases are staged once as two packed 2-bit streams (complemented little-endian for the
//     reverse complement, big-endian for the forward word) and every k-mer is CUT out of them with funnel shifts. Lane l
//     of a warp takes start 32 j + l, so the shift amounts are per-lane constants, the word indices advance by two per
//     step (immediate offsets), and the 8-byte stores of a warp are one contiguous 256-byte line -- no transposition;
//   * validity is a bit test on one bad-base bit row (and a sequence-start row for multi-sequence input): a k-mer is
//     valid iff its k bits (k-1 for the start row) are clear. Lane j of a warp computes the valid mask of the 32 starts
//     of step j in one go (log2 k doubling steps on a 96-bit window); tiles without any flagged base skip all of it and
//     write base + start index with unconditional stores;
//   * a CTA is 31 output warps + one SCAN warp. Tile i is staged (from registers requested one iteration earlier)
//     while tile i-1 is written out; the scan warp takes the tickets, runs the decoupled look-back of tile i and marks
//     the sequence starts of tile i+1 during that time, so nobody waits for the chained scan. Counts of flagged tiles
//     reach the scan warp through a named barrier the output warps only ARRIVE at; one CTA barrier per phase is left,
//     and tiles are large (27 776 starts; 7 936 for inputs below 20 Mbp) to amortise it.
//
// Measured (profiles/r01_s29_kmer2_kernel_details.txt, tools/kmerbench.py, 100 Mbp, value + position per k-mer):
// k=31 0.317 ms = 5.36 TB/s = 82 % of the measured HBM copy peak (first generation 0.446 ms), k=63 0.444 ms = 86 %
// (0.734 ms), 150-bp reads 0.43 ms (0.57), 1 % N 0.34 ms (0.44); 40 executed instructions per k-mer.
//
// SYNC = true adds a sampling predicate to the counting pass (k <= 32), so only the sampled k-mers are counted and written:
//   sync_on = 1  sampler::syncmer_sampler over hash::minimizer_position_extractor (include/syncmer_sampler.hpp:130-137,
//                include/kmer_view.hpp:266-283; blg_syncmers): every m-mer of a warp's span hashed once per strand into
//                shared-memory rows, each k-mer takes the minimum of its window; own tile size (K2S_SYNC steps).
//                1.64 ms per 100 Mbp at k=31 m=19 canonical (profiles/r01_s30_syncmer_bench_v3.txt)
//   sync_on = 2  sampler::hash_sampler (include/hash_sampler.hpp:138-141; blg_hash_sampled_kmers): hash64(value, seed) <
//                threshold, on the ordinary tiles. 0.62 ms per 100 Mbp (profiles/r01_s30_hash_sampler_bench.txt)
// Staging of the next tile's ASCII bytes: one bulk async copy per tile (cp.async.bulk global -> shared, completion on an mbarrier:
// UBLKCP in SASS) instead of one LDG.128 per thread held in registers across the write-out of the previous tile. Measured on
// B200, k=31 canonical, 100 Mbp: 0.3174 -> 0.3108 ms (81.8 % -> 83.5 % of the measured HBM copy peak), bit-identical
// (profiles/r02_r_kmer2_tma_staging.txt). -DK2_NO_TMA restores the register prefetch.
#ifndef K2_NO_TMA
#define K2_TMA 1
#endif
#include "common.cuh"
#include "kernels.h"

namespace blg {

namespace {
#ifndef K2_MIN_BLOCKS
#define K2_MIN_BLOCKS 1
#endif
#ifndef K2_SYNC_STEPS
#define K2_SYNC_STEPS 11
#endif
#ifndef K2_THREADS
#define K2_THREADS 992
#endif
constexpr int K2T = K2_THREADS;     // threads
constexpr int K2W = K2T / 32;       // warps
constexpr int K2PAD = 4;            // words in front of the big-endian stream

// S = steps (<= 32: lane j owns the valid mask of step j): a warp owns 32 S consecutive starts, a tile K2T * S.
// Two sizes are compiled: large tiles amortise the per-tile phases (barriers, look-back) on large inputs, small tiles
// keep every SM busy on small ones.
constexpr int K2S_LARGE = 28, K2S_SMALL = 8;
// the syncmer filter (k <= 32) keeps two hash rows per warp in shared memory: measured at 100 Mbp, k=31 m=19 canonical: 4 steps (two CTAs per SM) 2.33 ms, 8 steps 1.77 ms; 11 is the most that fits 227 KB
constexpr int K2S_SYNC = K2_SYNC_STEPS;
// non-canonical input needs the forward row only: twice the tile
constexpr int K2S_SYNC_FWD = 2 * K2_SYNC_STEPS;
template <int S>
struct K2Geo {
    static constexpr int TILE = K2T * S;        // starts per tile
    static constexpr int NW = TILE / 16 + 6;    // staged 16-base words: 15 (alignment) + TILE + 63 bases
    static constexpr int WORDS = NW + 8;        // every stream is padded so that cuts near the end stay inside
    static constexpr int ROW = WORDS / 2 + 2;
};

template <int S>
struct K2Buf {                       // one staged tile; two sets: tile i is staged while tile i-1 is written out
    u32 le[K2Geo<S>::WORDS];         // ~codes, base q of word j at bits [2q, 2q+2)
    u32 be[K2PAD + K2Geo<S>::WORDS]; // codes, first base of a word in its top bits
    u32 bad[K2Geo<S>::ROW];          // bit x: base A0 + x is not ACGTU (or outside the buffer)
    u32 seq[K2Geo<S>::ROW];          // bit x: a sequence starts at A0 + x (interior starts only)
    u32 bal[K2W][S];                 // valid masks of every warp's steps (tiles with flagged bases only)
};
template <int S>
struct K2Smem {
    K2Buf<S> buf[2];
    u32 wsum[K2W];
    u64 base[2];
    int next;
    int seqdirty[2];
#ifdef K2_TMA
    // staging experiment: the next tile's ASCII bytes arrive by ONE bulk async copy (cp.async.bulk,
    // completion on an mbarrier) instead of one LDG.128 per thread held in registers across the write-out of the previous tile
    alignas(16) unsigned char raw[(S == K2S_SYNC || S == K2S_SYNC_FWD) ? 16 : K2Geo<S>::NW * 16];   // the syncmer tiles keep the register prefetch
    alignas(8) unsigned long long mbar;
#endif
};


__device__ __forceinline__ u32 pairrev32_k(u32 x)
{
    x = __brev(x);
    return ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
}

// For the 32 consecutive starts x = bit .. bit + 31: bit x - bit of the result = OR of row bits [x, x + span), by ONE lane
// (log2(span) doubling steps on a 96-bit window instead of one span test per start). span <= 64.
__device__ __forceinline__ u32 or_span32(const u32* row, int bit, int span)
{
    if (span <= 0) return 0u;
    const int wq = bit >> 5, sh = bit & 31;
    const u32 w0 = row[wq], w1 = row[wq + 1], w2 = row[wq + 2], w3 = row[wq + 3];
    u32 x0 = __funnelshift_r(w0, w1, sh), x1 = __funnelshift_r(w1, w2, sh), x2 = __funnelshift_r(w2, w3, sh);
    int covered = 1;
    while (covered < span) {
        int step = covered < span - covered ? covered : span - covered;
        covered += step;
        u32 y0 = x0, y1 = x1, y2 = x2;
        if (step >= 32) { y0 = y1; y1 = y2; y2 = 0; step -= 32; }
        if (step) { y0 = __funnelshift_r(y0, y1, step); y1 = __funnelshift_r(y1, y2, step); y2 >>= step; }
        x0 |= y0; x1 |= y1; x2 |= y2;
    }
    return x0;
}

// Sequence-start bit row of the tile at `tile_lo` (interior starts inside the staged range), by one warp.
__device__ __forceinline__ int mark_tile_sequence_starts(const KmerArgs& a, long long tile_lo, u32* row, int nrow, int nw, int lane)
{
    for (int j = lane; j < nrow; j += 32) row[j] = 0;
    __syncwarp();
    const long long A0 = tile_lo & ~15LL;
    const long long lo_pos = A0 < 0 ? 0 : A0, end_pos = A0 + 16LL * nw;
    const long long lo = first_offset_at_or_after(a.seq_offsets, a.n_seq, lo_pos, a.buf_hi, lane);
    int found = 0;
    for (long long q = lo + lane; q < a.n_seq; q += 32) {
        const long long o = (long long)a.seq_offsets[q];
        if (o >= end_pos) break;
        const int rel = (int)(o - A0);
        atomicOr(&row[rel >> 5], 1u << (rel & 31));
        found = 1;
    }
    return __any_sync(0xffffffffu, found);
}
}  // namespace

// One step of the output pass: the k-mer at (per-lane constant shifts, word indices advanced by two per step) is cut
// out of the two streams; returns the value words.
template <bool WIDE, bool CANON>
__device__ __forceinline__ void cut_kmer(const u32* le, const u32* be, int s, int fs, const u32 (&vm)[4], u64& lo, u64& hi)
{
    constexpr int NV = WIDE ? 4 : 2;                            // 32-bit words of a value
    u32 r[NV], f[NV], c[NV + 1], d[NV + 1];
#pragma unroll
    for (int q = 0; q <= NV; ++q) { c[q] = le[q]; d[q] = be[q]; }
#pragma unroll
    for (int q = 0; q < NV; ++q) {
        r[q] = __funnelshift_r(c[q], c[q + 1], s) & vm[q];                          // reverse complement, low word first
        f[NV - 1 - q] = __funnelshift_l(d[q + 1], d[q], fs) & vm[NV - 1 - q];       // forward, d[0] holds the top bits
    }
    lo = ((u64)f[1] << 32) | f[0];
    hi = 0;
    if (WIDE) hi = ((u64)f[3] << 32) | f[2];
    if (CANON) {
        const u64 rlo = ((u64)r[1] << 32) | r[0];
        if (WIDE) {
            const u64 rhi = ((u64)r[3] << 32) | r[2];
            const bool lt = rhi < hi || (rhi == hi && rlo < lo);
            lo = lt ? rlo : lo; hi = lt ? rhi : hi;
        } else lo = rlo < lo ? rlo : lo;
    }
}

// Forward and reverse-complement value of the `len` (<= 32) bases at tile-local base index lp (mm = masks of 2 len bits).
__device__ __forceinline__ void cut_both(const u32* le_row, const u32* be_row, int lp, int len, const u32* mm, u64& f, u64& r)
{
    const int s = (lp & 15) * 2;
    const int e = 2 * lp + 2 * len + 32 * K2PAD - 64;             // the 64 bits that end with the last base
    const int fs = e & 31;
    const u32* le = le_row + (lp >> 4);
    const u32* be = be_row + (e >> 5);
    const u32 c0 = le[0], c1 = le[1], c2 = le[2], d0 = be[0], d1 = be[1], d2 = be[2];
    r = ((u64)(__funnelshift_r(c1, c2, s) & mm[1]) << 32) | (__funnelshift_r(c0, c1, s) & mm[0]);
    f = ((u64)(__funnelshift_l(d1, d0, fs) & mm[1]) << 32) | (__funnelshift_l(d2, d1, fs) & mm[0]);
}

// Syncmer filter only: per output warp, the hashes of its forward and reverse-complement m-mers (behind K2Smem).
template <int S>
struct K2SyncRow { static constexpr int N = 32 * S + 32; };
template <int S>
__device__ __forceinline__ u64* k2_sync_rows(unsigned char* raw)
{
    return reinterpret_cast<u64*>(raw + ((sizeof(K2Smem<S>) + 15) & ~(size_t)15));
}
template <int S>
constexpr size_t k2_smem_bytes(bool sync, bool both_strands = true)
{
    return ((sizeof(K2Smem<S>) + 15) & ~(size_t)15) + (sync ? (size_t)K2W * (both_strands ? 2 : 1) * K2SyncRow<S>::N * 8 : 0);
}

template <int S, bool WIDE, bool CANON, bool SYNC>
__global__ void __launch_bounds__(K2T + 32, K2_MIN_BLOCKS) kmer2_kernel(const KmerArgs a)
{
    constexpr int K2S = S, K2TILE = K2Geo<S>::TILE, K2NW = K2Geo<S>::NW, K2WORDS = K2Geo<S>::WORDS, K2ROW = K2Geo<S>::ROW;
    using K2Buf = K2Buf<S>;
    extern __shared__ __align__(16) unsigned char k2_raw[];
    K2Smem<S>& sm = *reinterpret_cast<K2Smem<S>*>(k2_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool is_scan = warp == K2W;                               // the extra warp: tickets and the look-back only
    const int k = a.k;
    // masks of the value words (k-mer = low 2k bits)
    u32 vm[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int bits = 2 * k - 32 * j;
        vm[j] = bits >= 32 ? 0xFFFFFFFFu : (bits > 0 ? ((1u << bits) - 1u) : 0u);
    }
    constexpr int NV = WIDE ? 4 : 2;
    const int wp = warp * (32 * K2S);                               // first start of this warp, tile-local

#ifdef K2_TMA
    if (tid == 0) mbar_init(&sm.mbar, 1);
    u32 tma_phase = 0;
#endif
    if (tid == 0) sm.next = (int)atomicAdd(a.tile_counter, 1u);
    __syncthreads();
    long long tile = (unsigned)sm.next;                             // the tile staged in this iteration
    __syncthreads();
    // sequence-start rows are the scan warp's: the row of a tile is marked one iteration before the tile is staged
    if (is_scan && a.seq_offsets != nullptr && tile < a.num_tiles) {
        const int f = mark_tile_sequence_starts(a, a.win_lo + tile * (long long)K2TILE, sm.buf[0].seq, K2ROW, K2NW, lane);
        if (lane == 0) sm.seqdirty[0] = f;
    }
    // the tile staged in the previous iteration, written out in this one (buffer set b ^ 1)
    bool have_prev = false, prev_all = false;
    long long prev_lo = 0;
    u32 prev_off = 0;
    int b = 0;
    // bases of the tile about to be staged, requested one iteration ahead (word tid, and K2T + tid)
    bool have_pre = false;
    uint4 pre0 = make_uint4(0, 0, 0, 0), pre1 = pre0;
    (void)pre1;                         // only the tiles with more than K2T staged words use the second register
    for (;;) {
        const bool have = tile < a.num_tiles;
        if (!have && !have_prev) break;
        if (have_prev && !prev_all && !is_scan) {
            // counts of the previous tile's warps (written before the last barrier, rewritten after the next one)
            prev_off = 0;
#pragma unroll
            for (int q = 0; q < K2W; ++q) {
                const u32 v = sm.wsum[q];
                if (q < warp) prev_off += v;
            }
        }
        K2Buf& cur = sm.buf[b];
        const long long tile_lo = a.win_lo + tile * (long long)K2TILE;
        const long long A0 = tile_lo & ~15LL;
        const int off0 = (int)(tile_lo - A0);
        int anybad = 0;
        if (have) {
            // the ticket after this one: every thread knows it after the next barrier, one whole iteration before it is staged
            if (is_scan && lane == 0) sm.next = (int)atomicAdd(a.tile_counter, 1u);
            // ---- stage: ASCII -> two packed streams + bad bits --------------------------------------------------------
            if (!is_scan) {
#ifdef K2_TMA
                if (have_pre) { mbar_wait(&sm.mbar, tma_phase); tma_phase ^= 1u; }
#endif
#pragma unroll
                for (int it = 0; it < (K2WORDS + K2T - 1) / K2T; ++it) {
                    const int wi = tid + it * K2T;
                    if (wi >= K2WORDS) break;
                    u32 codes = 0, bad16 = 0xFFFFu;
                    if (wi < K2NW) {
                        const long long P = A0 + 16LL * wi;
                        u32 w0, w1, w2, w3;
#ifdef K2_TMA
                        if (have_pre) {
                            const uint4 v = reinterpret_cast<const uint4*>(sm.raw)[wi];
                            w0 = v.x; w1 = v.y; w2 = v.z; w3 = v.w;
                        } else if
#else
                        if (have_pre && it < 2) {
                            const uint4 v = it == 0 ? pre0 : pre1;
                            w0 = v.x; w1 = v.y; w2 = v.z; w3 = v.w;
                        } else if
#endif
 (a.vec_ok && P >= a.buf_lo && P + 16 <= a.buf_hi) {
                            const uint4 v = __ldg(reinterpret_cast<const uint4*>(a.bases + (P - a.buf_lo)));
                            w0 = v.x; w1 = v.y; w2 = v.z; w3 = v.w;
                        } else {
                            u32 ww[4] = {0, 0, 0, 0};
#pragma unroll
                            for (int q = 0; q < 16; ++q) {
                                const long long pos = P + q;
                                if (pos >= a.buf_lo && pos < a.buf_hi) ww[q >> 2] |= (u32)a.bases[pos - a.buf_lo] << (8 * (q & 3));
                            }
                            w0 = ww[0]; w1 = ww[1]; w2 = ww[2]; w3 = ww[3];
                        }
                        blg_pack16_bits(w0, w1, w2, w3, &codes, &bad16);
                        anybad |= (int)bad16;
                    }
                    cur.le[wi] = ~codes;
                    cur.be[K2PAD + wi] = pairrev32_k(codes);
                    reinterpret_cast<unsigned short*>(cur.bad)[wi] = (unsigned short)bad16;
                }
                if (tid < K2PAD) cur.be[tid] = 0;
            }
        }
        const int dirty = __syncthreads_or(anybad);
        const long long next_tile = have ? (long long)(unsigned)sm.next : a.num_tiles;
        const bool seqdirty = have && a.seq_offsets != nullptr && sm.seqdirty[b];
        // every start of the tile is a valid k-mer inside the window range: no masks, no counting
        const bool all_valid = !SYNC && have && !dirty && !seqdirty && tile_lo + K2TILE <= a.win_hi;
        if (have && !all_valid) {
            // ---- valid masks of every warp's steps (kept in shared memory for the output pass), and their count -------
            if (!is_scan) {
                long long nwin = a.win_hi - (tile_lo + wp);     // starts of this warp inside the window range
                nwin = nwin < 0 ? 0 : (nwin > 32 * K2S ? 32 * K2S : nwin);
                // lane j owns step j: the valid mask of its 32 starts in one go
                u32 mrange = 0;
                if (lane < K2S) {
                    const int n = (int)nwin - 32 * lane;
                    mrange = n >= 32 ? 0xFFFFFFFFu : (n > 0 ? ((1u << n) - 1u) : 0u);
                    const int lp = off0 + wp + 32 * lane;
                    u32 bad = 0;
                    if (dirty) bad = or_span32(cur.bad, lp, k);
                    if (seqdirty) bad |= or_span32(cur.seq, lp + 1, k - 1);
                    mrange &= ~bad;
                    cur.bal[warp][lane] = mrange;
                }
                if (SYNC) {
                    // closed syncmers (include/syncmer_sampler.hpp:130-137 over include/kmer_view.hpp:266-283): of the valid
                    // k-mers keep those whose minimum-hash m-mer (hash64, seed 0; the extractor scans the bit fields of the
                    // k-mer VALUE from the right with '>=', so the leftmost field wins) sits at offset sync_s or sync_e.
                    // The reference hashes k-m+1 fields per k-mer; here every m-mer of the warp's span is hashed ONCE per
                    // strand into shared memory and a k-mer takes the minimum of its window: the fields of a forward value
                    // at start p are the forward m-mers at p .. p+w-1 (offset = position - p), those of a reverse-complement
                    // value are the reverse-complement m-mers of the same positions in reverse order (offset = p+w-1 -
                    // position: its leftmost field is the LAST position).
                    const int m = a.sync_m, w = k - m + 1;
                    u64* hf = k2_sync_rows<S>(k2_raw) + warp * ((CANON ? 2 : 1) * K2SyncRow<S>::N);
                    u64* hr = hf + K2SyncRow<S>::N;
                    if (a.sync_on == 2) {
                        // sampler::hash_sampler (include/hash_sampler.hpp:138-141): a k-mer is skipped while
                        // hash(value, seed) >= threshold
                        if (__ballot_sync(0xffffffffu, mrange != 0)) {
                            const int lp0 = off0 + wp + lane;
                            const int s = (lp0 & 15) * 2;
                            const int e = 2 * lp0 + 2 * k + 32 * K2PAD - 32 * NV;
                            const u32* le = cur.le + (lp0 >> 4);
                            const u32* be = cur.be + (e >> 5);
                            for (int j = 0; j < K2S; ++j) {
                                const u32 mj = __shfl_sync(0xffffffffu, mrange, j);
                                if (mj == 0) continue;
                                u64 km, hi_;
                                cut_kmer<WIDE, CANON>(le + 2 * j, be + 2 * j, s, e & 31, vm, km, hi_);
                                const u32 keep = __ballot_sync(0xffffffffu, murmur3_key8(km, a.samp_s8) < a.samp_threshold);
                                if (lane == j) mrange = mj & keep;
                            }
                        }
                    } else if (__ballot_sync(0xffffffffu, mrange != 0)) {
                        u32 mm[2];
                        mm[0] = 2 * m >= 32 ? 0xFFFFFFFFu : ((1u << (2 * m)) - 1u);
                        mm[1] = 2 * m >= 64 ? 0xFFFFFFFFu : (2 * m > 32 ? ((1u << (2 * m - 32)) - 1u) : 0u);
                        const int lpw = off0 + wp + lane;
                        for (int j = 0; j <= K2S; ++j) {
                            if (j == K2S && lane >= w - 1) break;
                            u64 f, r;
                            cut_both(cur.le, cur.be, lpw + 32 * j, m, mm, f, r);
                            hf[32 * j + lane] = murmur3_key8(f, 8ULL);              // seed 0: s8 = 0 ^ 8
                            if (CANON) hr[32 * j + lane] = murmur3_key8(r, 8ULL);
                        }
                        __syncwarp();
                        for (int j = 0; j < K2S; ++j) {
                            const u32 mj = __shfl_sync(0xffffffffu, mrange, j);
                            if (mj == 0) continue;
                            bool use_r = false;
                            if (CANON) {
                                u64 f, r;
                                cut_both(cur.le, cur.be, lpw + 32 * j, k, vm, f, r);
                                use_r = r < f;
                            }
                            // walk the fields in the order of the value's own string (offset 0 first): a strict '<' then
                            // keeps the leftmost minimum on either strand
                            const u64* h = use_r ? hr + 32 * j + lane + (w - 1) : hf + 32 * j + lane;
                            const int dir = use_r ? -1 : 1;
                            u64 best = h[0];
                            int off = 0;
#pragma unroll 4
                            for (int i = 1; i < w; ++i) {
                                const u64 v = h[i * dir];
                                if (v < best) { best = v; off = i; }
                            }
                            const u32 keep = __ballot_sync(0xffffffffu, off == a.sync_s || off == a.sync_e);
                            if (lane == j) mrange = mj & keep;
                        }
                        __syncwarp();
                    }
                    if (lane < K2S) cur.bal[warp][lane] = mrange;
                }
                u32 cnt = __popc(mrange);
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
                if (lane == 0) sm.wsum[warp] = cnt;
                // hand the counts to the scan warp without waiting for it (producer side of named barrier 1)
                __threadfence_block();
                asm volatile("bar.arrive 1, %0;" ::"r"(K2T + 32) : "memory");
            }
        }
        if (is_scan) {
            if (have) {
                u32 tile_total = K2TILE;
                if (!all_valid) {
                    asm volatile("bar.sync 1, %0;" ::"r"(K2T + 32) : "memory");
                    tile_total = 0;
#pragma unroll
                    for (int q = 0; q < K2W; ++q) tile_total += sm.wsum[q];
                }
                const u64 excl = lookback_exclusive(a.tile_status, tile, tile_total, lane);
                if (lane == 0) {
                    sm.base[b] = excl;
                    if (tile == a.num_tiles - 1) *a.total = excl + tile_total;
                }
            }
            if (a.seq_offsets != nullptr && next_tile < a.num_tiles) {
                // buffer set b ^ 1: its start row was last read by the counting pass of the previous iteration
                const int f = mark_tile_sequence_starts(a, a.win_lo + next_tile * (long long)K2TILE, sm.buf[b ^ 1].seq, K2ROW, K2NW, lane);
                if (lane == 0) sm.seqdirty[b ^ 1] = f;
            }
        } else {
            // ---- the next tile's bases are requested before the previous tile is written out ---------------------------
            {
                const long long nlo = a.win_lo + next_tile * (long long)K2TILE;
                const long long P0 = (nlo & ~15LL) - a.buf_lo;
                have_pre = next_tile < a.num_tiles && a.vec_ok && P0 >= 0 && P0 + 16LL * K2NW <= a.buf_hi - a.buf_lo;
#ifdef K2_TMA
                if (SYNC) have_pre = false;
#endif
#ifdef K2_TMA
                if (have_pre && tid == 0) bulk_load(sm.raw, a.bases + P0, (u32)(16 * K2NW), &sm.mbar);
#else
                if (have_pre) {
                    if (tid < K2NW) pre0 = __ldg(reinterpret_cast<const uint4*>(a.bases + P0) + tid);   // K2NW < K2T for the small tiles
                    if (K2T + tid < K2NW) pre1 = __ldg(reinterpret_cast<const uint4*>(a.bases + P0) + K2T + tid);
                }
#endif
            }
            // ---- write out the previous tile: its offset was resolved by the scan warp during the previous iteration ----
            if (have_prev) {
                const K2Buf& pv = sm.buf[b ^ 1];
                const int poff0 = (int)(prev_lo - (prev_lo & ~15LL));
                const int lp0 = poff0 + wp + lane;
                const int s = (lp0 & 15) * 2;
                const int e = 2 * lp0 + 2 * k + 32 * K2PAD - 32 * NV;   // the 32 NV bits that end with the k-mer's last base
                const int fs = e & 31;
                const u32* le = pv.le + (lp0 >> 4);
                const u32* be = pv.be + (e >> 5);
                const u64 g0 = sm.base[b ^ 1] + prev_off;
                const u64 p0 = (u64)(prev_lo + wp + lane);
                if (prev_all && g0 + 32 * K2S <= a.capacity) {
                    u64* vlo = a.value_lo + g0 + lane;
                    u64* vhi = WIDE ? a.value_hi + g0 + lane : nullptr;
                    u64* pos = a.position != nullptr ? a.position + g0 + lane : nullptr;
#pragma unroll
                    for (int j = 0; j < K2S; ++j) {
                        u64 lo, hi;
                        cut_kmer<WIDE, CANON>(le + 2 * j, be + 2 * j, s, fs, vm, lo, hi);
                        vlo[32 * j] = lo;
                        if (WIDE) vhi[32 * j] = hi;
                        if (pos != nullptr) pos[32 * j] = p0 + 32 * j;
                    }
                } else {
                    u64 g = g0;
                    const u32 below = (1u << lane) - 1u;
                    const bool fits = g0 + 32 * K2S <= a.capacity;      // this warp writes at most 32 K2S values
#pragma unroll 4
                    for (int j = 0; j < K2S; ++j) {
                        const u32 mj = prev_all ? 0xFFFFFFFFu : pv.bal[warp][j];
                        if (mj == 0) continue;
                        u64 lo, hi;
                        cut_kmer<WIDE, CANON>(le + 2 * j, be + 2 * j, s, fs, vm, lo, hi);
                        if (mj == 0xFFFFFFFFu && fits) {
                            const u64 o = g + lane;
                            a.value_lo[o] = lo;
                            if (WIDE) a.value_hi[o] = hi;
                            if (a.position != nullptr) a.position[o] = p0 + 32 * j;
                        } else {
                            const u64 o = g + __popc(mj & below);
                            if (((mj >> lane) & 1u) && o < a.capacity) {
                                a.value_lo[o] = lo;
                                if (WIDE) a.value_hi[o] = hi;
                                if (a.position != nullptr) a.position[o] = p0 + 32 * j;
                            }
                        }
                        g += __popc(mj);
                    }
                }
            }
        }
        __syncthreads();    // base[b] and the next ticket are out; buffer set b ^ 1 may be restaged
        have_prev = have; prev_all = all_valid; prev_lo = tile_lo; prev_off = (u32)wp;
        tile = next_tile;
        b ^= 1;
    }
}

namespace {
template <int S, bool WIDE, bool CANON, bool SYNC>
cudaError_t k2_setup(int* occ, bool rows = SYNC)
{
    cudaError_t e = cudaFuncSetAttribute(kmer2_kernel<S, WIDE, CANON, SYNC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)k2_smem_bytes<S>(rows, CANON));
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, kmer2_kernel<S, WIDE, CANON, SYNC>, K2T + 32, k2_smem_bytes<S>(rows, CANON));
}
template <int S>
cudaError_t k2_setup_s(bool wide, bool canon, int* occ)
{
    return wide ? (canon ? k2_setup<S, true, true, false>(occ) : k2_setup<S, true, false, false>(occ))
                : (canon ? k2_setup<S, false, true, false>(occ) : k2_setup<S, false, false, false>(occ));
}
template <int S>
void k2_launch_s(bool wide, bool canon, const KmerArgs& args, const KmerLaunchPlan& plan, cudaStream_t stream)
{
    if (wide) {
        if (canon) kmer2_kernel<S, true, true, false><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
        else kmer2_kernel<S, true, false, false><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    } else {
        if (canon) kmer2_kernel<S, false, true, false><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
        else kmer2_kernel<S, false, false, false><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    }
}
// large tiles once every CTA of the persistent grid gets five of them (tickets balance the rest)
bool k2_large(long long nwin, int sm_count) { return nwin >= 5LL * sm_count * K2Geo<K2S_LARGE>::TILE; }
}  // namespace

cudaError_t kmer2_plan(KmerArgs& args, int sm_count, KmerLaunchPlan* plan)
{
    const long long nwin = args.win_hi > args.win_lo ? args.win_hi - args.win_lo : 0;
    if (args.sync_on && args.k > 32) return cudaErrorInvalidValue;
    if (args.sync_on == 1 && (args.sync_m < 1 || args.sync_m > args.k)) return cudaErrorInvalidValue;
    // the syncmer filter has its own tile size (hash rows in shared memory); the hash-threshold filter needs no rows and
    // runs on the ordinary tiles
    const bool sync = args.sync_on == 1, samp = args.sync_on == 2;
    const bool large = !sync && k2_large(nwin, sm_count);
    const bool canon = args.canonical != 0, wide = args.k > 32;
    const bool sync_fwd = sync && !canon;
    const int tile = sync ? (sync_fwd ? K2Geo<K2S_SYNC_FWD>::TILE : K2Geo<K2S_SYNC>::TILE) : (large ? K2Geo<K2S_LARGE>::TILE : K2Geo<K2S_SMALL>::TILE);
    args.nwords = sync ? (sync_fwd ? K2Geo<K2S_SYNC_FWD>::NW : K2Geo<K2S_SYNC>::NW) : (large ? K2Geo<K2S_LARGE>::NW : K2Geo<K2S_SMALL>::NW);
    args.steps = sync ? (sync_fwd ? K2S_SYNC_FWD : K2S_SYNC) : (large ? K2S_LARGE : K2S_SMALL);
    args.num_tiles = (nwin + tile - 1) / tile;
    plan->threads = K2T + 32;
    plan->smem_bytes = sync ? (sync_fwd ? k2_smem_bytes<K2S_SYNC_FWD>(true, false) : k2_smem_bytes<K2S_SYNC>(true, true))
                            : (large ? k2_smem_bytes<K2S_LARGE>(false) : k2_smem_bytes<K2S_SMALL>(false));
    int occ = 0;
    cudaError_t e;
    if (sync) e = canon ? k2_setup<K2S_SYNC, false, true, true>(&occ) : k2_setup<K2S_SYNC_FWD, false, false, true>(&occ);
    else if (samp && large) e = canon ? k2_setup<K2S_LARGE, false, true, true>(&occ, false) : k2_setup<K2S_LARGE, false, false, true>(&occ, false);
    else if (samp) e = canon ? k2_setup<K2S_SMALL, false, true, true>(&occ, false) : k2_setup<K2S_SMALL, false, false, true>(&occ, false);
    else e = large ? k2_setup_s<K2S_LARGE>(wide, canon, &occ) : k2_setup_s<K2S_SMALL>(wide, canon, &occ);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
    long long grid = (long long)sm_count * occ;
    if (grid > args.num_tiles) grid = args.num_tiles;
    plan->grid = (int)(grid < 1 ? 1 : grid);
    return cudaSuccess;
}

cudaError_t kmer2_launch(const KmerArgs& args, const KmerLaunchPlan& plan, cudaStream_t stream)
{
    const bool canon = args.canonical != 0, wide = args.k > 32;
    if (args.sync_on == 1) {
        if (canon) kmer2_kernel<K2S_SYNC, false, true, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
        else kmer2_kernel<K2S_SYNC_FWD, false, false, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    } else if (args.sync_on == 2 && args.steps == K2S_LARGE) {
        if (canon) kmer2_kernel<K2S_LARGE, false, true, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
        else kmer2_kernel<K2S_LARGE, false, false, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    } else if (args.sync_on == 2) {
        if (canon) kmer2_kernel<K2S_SMALL, false, true, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
        else kmer2_kernel<K2S_SMALL, false, false, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    } else if (args.steps == K2S_LARGE) k2_launch_s<K2S_LARGE>(wide, canon, args, plan, stream);
    else k2_launch_s<K2S_SMALL>(wide, canon, args, plan, stream);
    return cudaGetLastError();
}

}  // namespace blg

// ---- blg_kmers_static_scratch: the k-mer kernel for callers that may not allocate ----------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc (code under
// CUDA graph capture, tight serving loops, harnesses that forbid it). Same kernel and same results as blg_kmers with device input
// and device output; its scan scratch (ticket counter, record total, one look-back word per tile) lives in the file-scope device
// array below instead of the context's cudaMalloc'ed buffers, so the call allocates nothing: no context, no cudaMalloc / cudaFree,
// no allocating library routine. Every GPU that loads the library holds the array (129 KB) until the process exits.
#include <cstdio>
#include <mutex>
#include "../../include/biolib_gpu.h"

namespace {

constexpr size_t K2_STATIC_MAX_BASES = (size_t)1 << 26;              // 64 Mi bases per call
constexpr int K2_STATIC_MAX_TILES = 16384;                           // look-back words (64 Mi bases are 8457 of the small tiles at most)
constexpr int K2_STATIC_CTRL = 32;                                   // words in front of them: [0] ticket counter, [1] record total
// In host code this name is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress (k2_static_setup).
static __device__ __align__(256) unsigned long long k2_static_scratch[K2_STATIC_CTRL + K2_STATIC_MAX_TILES];

struct K2StaticDev { unsigned long long* scratch; int sm_count; bool ready; };
K2StaticDev k2_static_dev[64] = {};
std::mutex k2_static_mutex;

// once per device, with that device current: the array's address on THIS device, and the device's SM count
cudaError_t k2_static_setup(int dev, K2StaticDev* out)
{
    std::lock_guard<std::mutex> lock(k2_static_mutex);
    if (!k2_static_dev[dev].ready) {
        unsigned long long* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), k2_static_scratch);
        if (e != cudaSuccess) return e;
        int sms = 0;
        if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
        k2_static_dev[dev] = {p, sms, true};
    }
    *out = k2_static_dev[dev];
    return cudaSuccess;
}

int k2_static_fail(char* err, size_t err_len, int status, const char* what, const char* detail = "")
{
    if (err && err_len) snprintf(err, err_len, "blg_kmers_static_scratch: %s%s", what, detail);
    return status;
}

}  // namespace

extern "C" int blg_kmers_static_scratch(const char* d_bases, size_t n_bases, uint8_t k, int canonical, uint64_t* d_value_lo, uint64_t* d_value_hi,
                                        uint64_t* d_position, uint64_t capacity, void* stream, uint64_t* d_count, uint64_t* count, char* err,
                                        size_t err_len)
{
    using namespace blg;
    if (err && err_len) err[0] = 0;
    if (count) *count = 0;
    if (k == 0 || k > 64) return k2_static_fail(err, err_len, BLG_E_INVALID, "need 1 <= k <= 64");
    if (n_bases > K2_STATIC_MAX_BASES)
        return k2_static_fail(err, err_len, BLG_E_CAPACITY, "n_bases exceeds the static scratch's limit of 67108864 (2^26) bases per call; use blg_kmers");
    if (n_bases && !d_bases) return k2_static_fail(err, err_len, BLG_E_INVALID, "d_bases is null");
    if (capacity && (!d_value_lo || (k > 32 && !d_value_hi))) return k2_static_fail(err, err_len, BLG_E_INVALID, "value arrays missing");
    if (!d_count && !count) return k2_static_fail(err, err_len, BLG_E_INVALID, "give d_count (device) or count (host; the call then waits for the stream)");
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return k2_static_fail(err, err_len, BLG_E_CUDA, "cudaGetDevice: ", cudaGetErrorString(e));
    if (dev < 0 || dev >= 64) return k2_static_fail(err, err_len, BLG_E_UNSUPPORTED, "device ordinal above 63");
    K2StaticDev sd;
    if ((e = k2_static_setup(dev, &sd)) != cudaSuccess) return k2_static_fail(err, err_len, BLG_E_CUDA, "cudaGetSymbolAddress: ", cudaGetErrorString(e));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const uint64_t nwin = n_bases >= (size_t)k ? n_bases - k + 1 : 0;

    KmerArgs a{};
    a.bases = reinterpret_cast<const unsigned char*>(d_bases);
    a.buf_lo = 0; a.buf_hi = (long long)n_bases;
    a.win_lo = 0; a.win_hi = (long long)nwin;
    a.k = k; a.canonical = canonical ? 1 : 0;
    a.vec_ok = (reinterpret_cast<uintptr_t>(d_bases) & 15u) == 0;
    a.value_lo = reinterpret_cast<unsigned long long*>(d_value_lo);
    a.value_hi = k > 32 ? reinterpret_cast<unsigned long long*>(d_value_hi) : nullptr;
    a.position = reinterpret_cast<unsigned long long*>(d_position);
    a.capacity = capacity;
    KmerLaunchPlan plan{};
    if (nwin) {
        if ((e = kmer2_plan(a, sd.sm_count, &plan)) != cudaSuccess) return k2_static_fail(err, err_len, BLG_E_CUDA, "kmer2_plan: ", cudaGetErrorString(e));
        if (a.num_tiles > K2_STATIC_MAX_TILES) return k2_static_fail(err, err_len, BLG_E_CAPACITY, "more tiles than the static scratch holds");
    }
    // the array keeps the last call's contents: what the kernel expects to be zero is cleared here, on the caller's stream
    const size_t clear_words = (size_t)K2_STATIC_CTRL + (size_t)(nwin ? a.num_tiles : 0);
    if ((e = cudaMemsetAsync(sd.scratch, 0, clear_words * 8, st)) != cudaSuccess) return k2_static_fail(err, err_len, BLG_E_CUDA, "cudaMemsetAsync: ", cudaGetErrorString(e));
    if (nwin) {
        a.tile_counter = reinterpret_cast<unsigned int*>(sd.scratch);
        a.total = sd.scratch + 1;
        a.tile_status = sd.scratch + K2_STATIC_CTRL;
        if ((e = kmer2_launch(a, plan, st)) != cudaSuccess) return k2_static_fail(err, err_len, BLG_E_CUDA, "launch: ", cudaGetErrorString(e));
    }
    if (d_count && (e = cudaMemcpyAsync(d_count, sd.scratch + 1, 8, cudaMemcpyDeviceToDevice, st)) != cudaSuccess)
        return k2_static_fail(err, err_len, BLG_E_CUDA, "cudaMemcpyAsync: ", cudaGetErrorString(e));
    if (count) {
        unsigned long long total = 0;
        if ((e = cudaMemcpyAsync(&total, sd.scratch + 1, 8, cudaMemcpyDeviceToHost, st)) != cudaSuccess || (e = cudaStreamSynchronize(st)) != cudaSuccess)
            return k2_static_fail(err, err_len, BLG_E_CUDA, "reading the count: ", cudaGetErrorString(e));
        *count = total;
        if (total > capacity) return k2_static_fail(err, err_len, BLG_E_CAPACITY, "output capacity too small (*count holds the required size)");
    }
    return BLG_OK;
}
