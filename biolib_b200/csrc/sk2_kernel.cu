// sk2_kernel.cu -- fused minimizer / super-k-mer extraction for sm_100a, second generation.
//
// Same results as sk_kernel.cu (SURVEY.md 8(a) decisions D1-D8; reference intent include/minimizer_view.hpp:184-266,
// include/super_kmer_view.hpp:121-135), organised around the ALU pipe, which is what bounds this path on B200
// (profiles/: one MurmurHash3 per base is ~21 ALU + ~18 IMAD lane-ops before anything else happens):
//
//   * a thread owns 16 consecutive m-mer starts aligned to a 16-base word, so every funnel shift that pulls an
//     m-mer out of the packed streams has an immediate shift amount and nothing is rolled or warmed up;
//   * two packed streams are staged per tile: the complemented little-endian stream (reverse-complement word =
//     one funnel shift + mask) and a big-endian stream delayed by 66-2m bits (forward word = one funnel shift +
//     mask with the same immediates for every m);
//   * m-mer validity is never computed: an m-mer that touches a bad byte only occurs in windows (k-mers) that are
//     invalid, and k-mer validity is a bit-parallel OR-span over the tile's bad-byte bits, skipped for clean tiles;
//   * the window minimum is a sparse-table doubling (span 1, 2, 4, ... <= w, then one combine at offset
//     w - 2^L) over (hash, position) carried in registers, with only the first 2^(l-1) entries of each thread's
//     level going through shared memory: about log2(w)+1 compare/selects per window instead of three scans over
//     shared-memory arrays. Ties keep the left operand, so the leftmost minimum wins exactly as the strict '<'
//     of include/minimizer_view.hpp:283,374 does;
//   * a record boundary is simply "argmin position differs from the next window's", one 32-bit compare;
//   * one WARP owns a tile of 512 m-mer starts end to end (only __syncwarp inside), so the warps of an SM drift apart
//     and cover each other's latencies; the four warps of a CTA form one scan unit and meet at ONE barrier per tile;
//   * record starts are compacted with popc + warp scan; the ordered global offset comes from a decoupled look-back
//     over scan units that is software-pipelined: a unit publishes its count when it is known, resolves its prefix
//     half a tile later (when all predecessors have long published, so nobody polls) and writes its records one
//     tile later, from a second set of shared-memory buffers, record-parallel with coalesced stores.
#include "common.cuh"
#include "kernels.h"

namespace blg {

namespace {

constexpr int NT = 32;            // lanes: a warp owns a tile
constexpr int WARPS = SK2_THREADS / 32;
constexpr int Q = 16 * NT;        // m-mer starts hashed per tile
constexpr int NW = NT + 2;        // staged 16-base words
constexpr int ML_ENTRIES = Q / 16 * 17 + 40;   // 17-entry pitch per lane + slack for reads at +df, +16

// Shared memory of one warp, as byte offsets from the warp's base (plain offsets keep every access an LDS/STS).
// BUF0 / BUF1 are double buffered: the records of tile i are written out while tile i+1 is being computed.
constexpr int OFF_E0 = 0;                        // u32 [NW + 2]  codes, big-endian inside the word (base q at bits 2(15-q))
constexpr int OFF_BADW = OFF_E0 + 144;           // u32 [NW + 8]  16 bad-byte bits per word
constexpr int OFF_SEQW = OFF_BADW + 176;         // u32 [NW + 8]  16 sequence-start bits per word
constexpr int OFF_FLAG = OFF_SEQW + 176;         // int [4]       any-sequence-start flag
constexpr int OFF_BUF = OFF_FLAG + 16;
constexpr int BUF_CW = 0;                        // u32 [NW + 2]  complemented codes, little-endian (base q at bits 2q)
constexpr int BUF_FW = BUF_CW + 144;             // u32 [NW + 2]  big-endian stream delayed by 66-2m bits
constexpr int BUF_IDXF = BUF_FW + 144;           // u16 [Q + 16]  argmin position of every window
constexpr int BUF_REC = BUF_IDXF + (Q + 16) * 2; // u16 [Q]       record starts
constexpr int BUF_XM = BUF_REC + Q * 2;          // u32 [NT + 8]  low 16: "no continuation" bits, high 16: record-end bits
constexpr int BUF_BYTES = BUF_XM + (NT + 8) * 4;
constexpr int OFF_ML = OFF_BUF + 2 * BUF_BYTES;  // u32 [3][ML_ENTRIES] last doubling level (hi | lo | position);
                                                 // the partial levels (uint4) use the same bytes before it is written
constexpr int ML_BYTES = 3 * ML_ENTRIES * 4;
static_assert(OFF_ML % 16 == 0 && BUF_BYTES % 16 == 0 && OFF_BUF % 16 == 0, "16-byte aligned regions");

__host__ __device__ inline size_t sk2_warp_bytes(int pp_entries)
{
    const size_t pp = (size_t)pp_entries * 16;
    return (size_t)OFF_ML + (pp > (size_t)ML_BYTES ? pp : (size_t)ML_BYTES);
}

__device__ __forceinline__ u32 pairrev32(u32 x)
{
    x = __brev(x);
    return ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
}

// b replaces a only when strictly smaller: a always covers the positions to the left
__device__ __forceinline__ void keep_min(u32& ahi, u32& alo, u32& aix, u32 bhi, u32 blo, u32 bix)
{
    const u64 av = ((u64)ahi << 32) | alo, bv = ((u64)bhi << 32) | blo;
    const bool lt = bv < av;
    ahi = lt ? bhi : ahi;
    alo = lt ? blo : alo;
    aix = lt ? bix : aix;
}

// One doubling level: entry i := min(entry i, entry i + D). Entries >= 16 live in the lane(s) to the right and
// arrive through `pp` (first min(D,16) entries of every lane's previous level, lane pitch = that count).
template <int D>
__device__ __forceinline__ void doubling_level(u32 (&hi)[16], u32 (&lo)[16], u32 (&ix)[16], uint4* pp, int t)
{
    constexpr int NB = D < 16 ? D : 16;
    constexpr int TOFF = D <= 16 ? 1 : D / 16;
#pragma unroll
    for (int j = 0; j < NB; ++j) pp[t * NB + j] = make_uint4(hi[j], lo[j], ix[j], 0u);
    __syncwarp();
    uint4 nb[NB];
#pragma unroll
    for (int j = 0; j < NB; ++j) nb[j] = pp[(t + TOFF) * NB + j];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (D < 16 && i + D < 16) keep_min(hi[i], lo[i], ix[i], hi[i + D], lo[i + D], ix[i + D]);
        else {
            const uint4 e = nb[D < 16 ? i + D - 16 : i];
            keep_min(hi[i], lo[i], ix[i], e.x, e.y, e.z);
        }
    }
}

// low 17 bits: OR over d in [0, span) of bit (16 t + i + d + shift) of the per-word 16-bit rows, for i = 0..16
__device__ __forceinline__ u32 or_span17(const u32* rows, int t, int span, int shift)
{
    u32 x0 = rows[t] | (rows[t + 1] << 16), x1 = rows[t + 2] | (rows[t + 3] << 16), x2 = rows[t + 4] | (rows[t + 5] << 16);
    if (shift) { x0 = __funnelshift_r(x0, x1, 1); x1 = __funnelshift_r(x1, x2, 1); x2 >>= 1; }
    int covered = 1;
    while (covered < span) {
        int step = covered < span - covered ? covered : span - covered;
        covered += step;
        u32 y0 = x0, y1 = x1, y2 = x2;
        if (step >= 32) { y0 = y1; y1 = y2; y2 = 0; step -= 32; }
        if (step) { y0 = __funnelshift_r(y0, y1, step); y1 = __funnelshift_r(y1, y2, step); y2 >>= step; }
        x0 |= y0; x1 |= y1; x2 |= y2;
    }
    return x0 & 0x1FFFFu;
}

struct TileConst {
    u32 maskhi, masklo;
    int dl_words, dl_bits;
    int w, m, k, L, df;
    u64 s8;
};

// Everything of one tile up to the list of record starts, by one warp, into buffer set b. Returns the number of
// records that start in the tile. `resolve_unit >= 0`: this warp also resolves that (earlier) scan unit's prefix,
// after the window minima, where nothing but the argmin positions is live.
template <bool CANON, bool MW>
__device__ __forceinline__ u32 compute_tile(const SkArgs& a, unsigned char* wsm, const TileConst& c, int b, long long tile, int t,
                                            bool have_pre, uint4 pre0, uint4 pre1,
                                            long long resolve_unit, u64 resolve_agg, u64* resolve_out)
{
    const long long tbase = a.org + tile * (long long)a.R - 16;    // global position of local 0 (multiple of 16)
    unsigned char* buf = wsm + OFF_BUF + b * BUF_BYTES;
    u32* E0 = reinterpret_cast<u32*>(wsm + OFF_E0);
    u32* BADW = reinterpret_cast<u32*>(wsm + OFF_BADW);
    u32* SEQW = reinterpret_cast<u32*>(wsm + OFF_SEQW);
    int* FLAG = reinterpret_cast<int*>(wsm + OFF_FLAG);
    uint4* PP = reinterpret_cast<uint4*>(wsm + OFF_ML);
    u32* MLH = reinterpret_cast<u32*>(wsm + OFF_ML);
    u32* MLL = MLH + ML_ENTRIES;
    u32* MLX = MLL + ML_ENTRIES;
    u32* CW = reinterpret_cast<u32*>(buf + BUF_CW);
    u32* FW = reinterpret_cast<u32*>(buf + BUF_FW);
    unsigned short* IDXF = reinterpret_cast<unsigned short*>(buf + BUF_IDXF);
    unsigned short* REC = reinterpret_cast<unsigned short*>(buf + BUF_REC);
    u32* XM = reinterpret_cast<u32*>(buf + BUF_XM);

    // ---- stage: ASCII -> packed streams + bad-byte bits -----------------------------------------------------------
    int anybad = 0;
    for (int wi = t; wi < NW + 8; wi += NT) {
        u32 codes = 0, bad = 0xFFFFu;
        if (wi < NW) {
            const long long P = tbase + 16LL * wi;
            u32 w0, w1, w2, w3;
            if (have_pre) {
                const uint4 v = wi < NT ? pre0 : pre1;
                w0 = v.x; w1 = v.y; w2 = v.z; w3 = v.w;
            } else if (a.vec_ok && P >= a.buf_lo && P + 16 <= a.buf_hi) {
                const uint4 v = __ldg(reinterpret_cast<const uint4*>(a.bases + (P - a.buf_lo)));
                w0 = v.x; w1 = v.y; w2 = v.z; w3 = v.w;
            } else {
                w0 = w1 = w2 = w3 = 0;                 // tile edges / unaligned buffers: byte loads, rolled
#pragma unroll 1
                for (int q = 0; q < 16; ++q) {
                    const long long pos = P + q;
                    const u32 byte = (pos >= a.buf_lo && pos < a.buf_hi) ? (u32)a.bases[pos - a.buf_lo] : 0u;
                    const u32 sh = byte << (8 * (q & 3));
                    if (q < 4) w0 |= sh; else if (q < 8) w1 |= sh; else if (q < 12) w2 |= sh; else w3 |= sh;
                }
            }
            blg_pack16_bits(w0, w1, w2, w3, &codes, &bad);
            CW[wi] = ~codes;
            E0[wi] = pairrev32(codes);
            anybad |= (bad != 0);
        }
        BADW[wi] = bad;
        SEQW[wi] = 0;
    }
    if (t == 0) FLAG[0] = 0;
    __syncwarp();
    const int dirty = __any_sync(0xffffffffu, anybad);
    if (a.seq_offsets != nullptr) {
        // interior sequence starts inside the staged range
        long long lo = 1, hi = a.n_seq;
        const long long lo_pos = tbase < 0 ? 0 : tbase, end_pos = tbase + 16LL * NW;
        while (lo < hi) {
            const long long mid = (lo + hi) >> 1;
            if ((long long)a.seq_offsets[mid] < lo_pos) lo = mid + 1; else hi = mid;
        }
        int found = 0;
        for (long long q = lo + t; q < a.n_seq; q += 32) {
            const long long o = (long long)a.seq_offsets[q];
            if (o >= end_pos) break;
            const int rel = (int)(o - tbase);
            atomicOr(&SEQW[rel >> 4], 1u << (rel & 15));
            found = 1;
        }
        if (found) FLAG[0] = 1;
    }
    for (int x = t; x < NW + 2; x += NT) {
        const int i1 = x - c.dl_words, i0 = i1 - 1;
        const u32 lo_w = (i1 >= 0 && i1 < NW) ? E0[i1] : 0u, hi_w = (i0 >= 0 && i0 < NW) ? E0[i0] : 0u;
        FW[x] = __funnelshift_r(lo_w, hi_w, c.dl_bits);
    }
    __syncwarp();
    const int seqdirty = (a.seq_offsets != nullptr) && FLAG[0];

    // ---- one hash per m-mer start ---------------------------------------------------------------------------------
    // A loop of 4 x 4 instead of 16 unrolled copies keeps the kernel inside the 32 KB instruction cache (16 warps per
    // SM sit at different program counters). The hashes take a round trip through shared memory (level-0 array,
    // 17-entry pitch) because a rolled loop cannot index registers.
    {
        const u32 W0 = FW[t], W1 = FW[t + 1], W2 = FW[t + 2];
        const u32 C0 = CW[t], C1 = CW[t + 1], C2 = CW[t + 2];
        u32* hh = MLH + 17 * t;
        u32* hl = MLL + 17 * t;
#pragma unroll 1
        for (int it = 0; it < 4; ++it) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int i = 4 * it + j;
                u32 flo = __funnelshift_r(W2, W1, 30 - 2 * i), fhi = 0, rlo = 0, rhi = 0;
                if (MW) fhi = __funnelshift_r(W1, W0, 30 - 2 * i) & c.maskhi; else flo &= c.masklo;
                u64 val = ((u64)fhi << 32) | flo;
                if (CANON) {
                    rlo = __funnelshift_r(C0, C1, 2 * i);
                    if (MW) rhi = __funnelshift_r(C1, C2, 2 * i) & c.maskhi; else rlo &= c.masklo;
                    const u64 rc = ((u64)rhi << 32) | rlo;
                    val = rc < val ? rc : val;
                }
                const u64 h = murmur3_key8(val, c.s8);
                hh[i] = (u32)(h >> 32); hl[i] = (u32)h;
            }
        }
    }
    __syncwarp();
    u32 hi[16], lo[16], ix[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) { hi[i] = MLH[17 * t + i]; lo[i] = MLL[17 * t + i]; ix[i] = (u32)(16 * t + i); }

    // ---- window minimum by doubling ---------------------------------------------------------------------------------
    {
        const u32 nbh = MLH[17 * (t + 1)], nbl = MLL[17 * (t + 1)];
        __syncwarp();   // level 0 is in registers: the partial levels may now overwrite these bytes
        if (c.L >= 1) {
#pragma unroll
            for (int i = 0; i < 15; ++i) keep_min(hi[i], lo[i], ix[i], hi[i + 1], lo[i + 1], ix[i + 1]);
            keep_min(hi[15], lo[15], ix[15], nbh, nbl, (u32)(16 * t + 16));
        }
    }
    if (c.L >= 2) doubling_level<2>(hi, lo, ix, PP + a.pp_off[2], t);
    if (c.L >= 3) doubling_level<4>(hi, lo, ix, PP + a.pp_off[3], t);
    if (c.L >= 4) doubling_level<8>(hi, lo, ix, PP + a.pp_off[4], t);
    if (c.L >= 5) doubling_level<16>(hi, lo, ix, PP + a.pp_off[5], t);
    if (c.L >= 6) doubling_level<32>(hi, lo, ix, PP + a.pp_off[6], t);
    // last level in full, 17-entry pitch per lane; window s = min(level[s], level[s + df])
    __syncwarp();   // every lane is done reading the partial levels, which share these bytes
#pragma unroll
    for (int i = 0; i < 16; ++i) { MLH[17 * t + i] = hi[i]; MLL[17 * t + i] = lo[i]; MLX[17 * t + i] = ix[i]; }
    __syncwarp();
    u32 vix[17];
    {
        const int cross = 16 - (c.df & 15);           // first i whose partner sits in the next lane's row
        const int rowA = 17 * (t + (c.df >> 4)) + (c.df & 15);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const int e = rowA + i + (i >= cross ? 1 : 0);   // one pad entry per row
            const u64 av = ((u64)hi[i] << 32) | lo[i], bv = ((u64)MLH[e] << 32) | MLL[e];
            vix[i] = bv < av ? MLX[e] : ix[i];
        }
        const int l16 = 17 * (t + 1), r16 = rowA + 17;
        const u64 av = ((u64)MLH[l16] << 32) | MLL[l16], bv = ((u64)MLH[r16] << 32) | MLL[r16];
        vix[16] = bv < av ? MLX[r16] : MLX[l16];
    }

    if (resolve_unit >= 0) {
        const u64 excl = scan_resolve(a.tile_status, resolve_unit, resolve_agg, t);
        if (t == 0) *resolve_out = excl;
    }

    // ---- validity, boundaries -----------------------------------------------------------------------------------------
    u32 V = 0x1FFFFu;
    if (dirty | seqdirty) {
        u32 inv = 0;
        if (dirty) inv |= or_span17(BADW, t, c.k, 0);
        if (seqdirty && c.k > 1) inv |= or_span17(SEQW, t, c.k - 1, 1);
        V = ~inv & 0x1FFFFu;
    }
    u32 dif = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) dif |= (u32)(vix[i + 1] != vix[i]) << i;
    const u32 E = V & (~(V >> 1) | dif) & 0xFFFFu;   // the record ends with this window
    const u32 X = (E | ~V) & 0xFFFFu;                // the next window cannot continue a record
    XM[t] = X | (E << 16);
    {
        uint4 p0, p1;
        p0.x = vix[0] | (vix[1] << 16); p0.y = vix[2] | (vix[3] << 16); p0.z = vix[4] | (vix[5] << 16); p0.w = vix[6] | (vix[7] << 16);
        p1.x = vix[8] | (vix[9] << 16); p1.y = vix[10] | (vix[11] << 16); p1.z = vix[12] | (vix[13] << 16); p1.w = vix[14] | (vix[15] << 16);
        uint4* dst = reinterpret_cast<uint4*>(IDXF + 16 * t);
        dst[0] = p0; dst[1] = p1;
    }
    __syncwarp();

    // ---- record starts ---------------------------------------------------------------------------------------------------
    u32 start = 0;
    if (t >= 1 && t <= a.R / 16) {
        const u32 xprev = (XM[t - 1] >> 15) & 1u;
        start = V & ((X << 1) | xprev) & 0xFFFFu;
        // windows outside [win_lo, win_hi) belong to another call / rank
        const long long g0 = tbase + 16LL * t;
        long long nlo = a.win_lo - g0, nhi = a.win_hi - g0;
        nlo = nlo < 0 ? 0 : (nlo > 16 ? 16 : nlo);
        nhi = nhi < 0 ? 0 : (nhi > 16 ? 16 : nhi);
        start &= (0xFFFFu << nlo) & ((1u << nhi) - 1u);
    }
    const u32 cnt = __popc(start);
    u32 incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const u32 v = __shfl_up_sync(0xffffffffu, incl, o);
        if (t >= o) incl += v;
    }
    u32 off = incl - cnt, sm = start;
    while (sm) {
        const int i = __ffs((int)sm) - 1;
        sm &= sm - 1;
        REC[off++] = (unsigned short)(16 * t + i);
    }
    return __shfl_sync(0xffffffffu, incl, 31);
}

// Record-parallel output of a finished tile from buffer set b: `n_rec` records, the first one at global index `base`.
template <bool CANON, bool MW>
__device__ __forceinline__ void emit_tile(const SkArgs& a, const unsigned char* wsm, const TileConst& c, int b, long long tile, int t,
                                          u32 n_rec, u64 base)
{
    const long long tbase = a.org + tile * (long long)a.R - 16;
    const unsigned char* buf = wsm + OFF_BUF + b * BUF_BYTES;
    const u32* CW = reinterpret_cast<const u32*>(buf + BUF_CW);
    const u32* FW = reinterpret_cast<const u32*>(buf + BUF_FW);
    const unsigned short* IDXF = reinterpret_cast<const unsigned short*>(buf + BUF_IDXF);
    const unsigned short* REC = reinterpret_cast<const unsigned short*>(buf + BUF_REC);
    const u32* XM = reinterpret_cast<const u32*>(buf + BUF_XM);
    for (u32 r = t; r < n_rec; r += NT) {
        const int s0 = REC[r];
        const int j = IDXF[s0];
        // size: distance to the first record-end bit at or after s0, at most w
        int tw = s0 >> 4, consumed = 0, size = c.w;
        u32 cur = (XM[tw] >> 16) >> (s0 & 15);
        int avail = 16 - (s0 & 15);
        for (;;) {
            if (cur) { const int f = consumed + __ffs((int)cur); size = f < c.w ? f : c.w; break; }
            consumed += avail;
            if (consumed >= c.w) break;
            cur = XM[++tw] >> 16;
            avail = 16;
        }
        const int tj = j >> 4, ij = j & 15;
        u32 flo = __funnelshift_r(FW[tj + 2], FW[tj + 1], 30 - 2 * ij), fhi = 0;
        if (MW) fhi = __funnelshift_r(FW[tj + 1], FW[tj], 30 - 2 * ij) & c.maskhi; else flo &= c.masklo;
        u64 val = ((u64)fhi << 32) | flo;
        if (CANON) {
            u32 rlo = __funnelshift_r(CW[tj], CW[tj + 1], 2 * ij), rhi = 0;
            if (MW) rhi = __funnelshift_r(CW[tj + 1], CW[tj + 2], 2 * ij) & c.maskhi; else rlo &= c.masklo;
            const u64 rc = ((u64)rhi << 32) | rlo;
            val = rc < val ? rc : val;
        }
        const u64 g = base + r;
        if (g < a.capacity) {
            a.value[g] = val;
            store_mm_start(a, g, (long long)(tbase + j));
            if (a.size != nullptr) { a.size[g] = (u8)size; a.mm_pos[g] = (u8)(j - s0); }
            if (a.hash != nullptr) a.hash[g] = murmur3_key8(val, c.s8);
        }
    }
    __syncwarp();   // this buffer set is restaged by the next compute_tile
}

}  // namespace

template <bool CANON, bool MW>
__global__ void __launch_bounds__(SK2_THREADS, 512 / SK2_THREADS) sk2_kernel(const SkArgs a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int t = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char* wsm = smem_raw + warp * a.warp_smem;
    // CTA-wide words behind the per-warp regions
    u32* CNT = reinterpret_cast<u32*>(smem_raw + (size_t)WARPS * a.warp_smem);   // [2][WARPS] record counts
    u64* UBASE = reinterpret_cast<u64*>(CNT + 2 * WARPS);                        // [2] resolved prefix of a unit
    u32* NEXT = reinterpret_cast<u32*>(UBASE + 2);                               // [2] ticket of the following unit

    TileConst c;
    c.w = a.w; c.m = a.m; c.k = a.k; c.L = a.levels; c.df = a.df; c.s8 = a.s8;
    c.maskhi = MW ? (c.m == 32 ? 0xFFFFFFFFu : ((1u << (2 * c.m - 32)) - 1u)) : 0u;
    c.masklo = MW ? 0xFFFFFFFFu : (c.m == 16 ? 0xFFFFFFFFu : ((1u << (2 * c.m)) - 1u));
    const int delay = 66 - 2 * c.m;
    c.dl_words = delay >> 5; c.dl_bits = delay & 31;

    const long long num_units = (a.num_tiles + WARPS - 1) / WARPS;
    if (threadIdx.x == 0) NEXT[0] = atomicAdd(a.tile_counter, 1u);
    __syncthreads();

    // Pipeline over scan units (WARPS consecutive tiles, one per warp). In iteration i:
    //   every warp computes its tile of unit i into buffer set b;
    //   one warp (rotating) resolves the prefix of unit i-1 in the middle of that;
    //   barrier; unit i's aggregate is published; every warp writes out its tile of unit i-1 from buffer set b^1.
    long long unit = NEXT[0], prev_unit = -1;
    bool have_pre = false;
    uint4 pre0 = make_uint4(0, 0, 0, 0), pre1 = pre0;
    u32 prev_n = 0, prev_pre = 0;      // this warp's record count / count of the warps before it, in the previous unit
    u64 prev_agg = 0;                  // the previous unit's total (kept by every warp: the resolver rotates)
    int b = 0;
    for (int iter = 0;; ++iter) {
        const bool have = unit < num_units;
        if (!have && prev_unit < 0) break;
        if (threadIdx.x == 0) NEXT[b ^ 1] = have ? atomicAdd(a.tile_counter, 1u) : 0xFFFFFFFFu;
        const bool resolver = prev_unit >= 0 && warp == (iter % WARPS);
        const long long tile = unit * WARPS + warp;
        u32 n_rec = 0;
        if (have && tile < a.num_tiles)
            n_rec = compute_tile<CANON, MW>(a, wsm, c, b, tile, t, have_pre, pre0, pre1, resolver ? prev_unit : -1, prev_agg, &UBASE[b ^ 1]);
        else if (resolver) {
            const u64 excl = scan_resolve(a.tile_status, prev_unit, prev_agg, t);
            if (t == 0) UBASE[b ^ 1] = excl;
        }
        if (t == 0) CNT[b * WARPS + warp] = n_rec;
#ifndef X_NOBAR
        __syncthreads();
#else
        __syncwarp();
#endif
        u32 pre = 0, agg = 0;
#pragma unroll
        for (int q = 0; q < WARPS; ++q) {
            const u32 v = CNT[b * WARPS + q];
            if (q < warp) pre += v;
            agg += v;
        }
        if (have && threadIdx.x == 0) scan_publish(a.tile_status, unit, agg);
        if (prev_unit >= 0) {
            const u64 ubase = UBASE[b ^ 1];
            if (threadIdx.x == 0 && prev_unit == num_units - 1) *a.total = ubase + prev_agg;
        }
        const long long next_unit = NEXT[b ^ 1];
        // the next tile's bases start their trip from HBM now and arrive while the previous tile's records are written
        have_pre = false;
        {
            const long long nt = next_unit * WARPS + warp;
            const long long P0 = a.org + nt * (long long)a.R - 16 - a.buf_lo;
            if (next_unit < num_units && nt < a.num_tiles && a.vec_ok && P0 >= 0 && P0 + 16LL * NW <= a.buf_hi - a.buf_lo) {
                have_pre = true;
                pre0 = __ldg(reinterpret_cast<const uint4*>(a.bases + P0) + t);
                if (t < NW - NT) pre1 = __ldg(reinterpret_cast<const uint4*>(a.bases + P0) + NT + t);
            }
        }
        if (prev_unit >= 0)
            emit_tile<CANON, MW>(a, wsm, c, b ^ 1, prev_unit * WARPS + warp, t, prev_n, UBASE[b ^ 1] + prev_pre);
        prev_unit = have ? unit : -1;
        prev_n = n_rec; prev_pre = pre; prev_agg = agg;
        unit = next_unit;
        b ^= 1;
    }
}

template <bool CANON, bool MW>
static cudaError_t sk2_setup(size_t smem, int* occ)
{
    cudaError_t e = cudaFuncSetAttribute(sk2_kernel<CANON, MW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, sk2_kernel<CANON, MW>, SK2_THREADS, smem);
}

cudaError_t sk2_plan(SkArgs& args, int sm_count, SkLaunchPlan* plan)
{
    const int w = args.w;
    int L = 0;
    while ((2 << L) <= w) ++L;          // largest L with 2^L <= w
    args.levels = L;
    args.df = w - (1 << L);
    int entries = 0;
    for (int l = 0; l < 8; ++l) args.pp_off[l] = 0;
    for (int l = 2; l <= L; ++l) {
        const int d = 1 << (l - 1), nb = d < 16 ? d : 16;
        args.pp_off[l] = entries;
        entries += (NT + 2) * nb;
    }
    args.pp_entries = entries;
    args.R = (Q - 2 * w - 14) / 16 * 16;
    args.org = args.win_lo & ~15LL;
    const long long nwin = args.win_hi > args.org ? args.win_hi - args.org : 0;
    args.num_tiles = args.win_hi > args.win_lo ? (nwin + args.R - 1) / args.R : 0;
    plan->threads = SK2_THREADS;
    args.warp_smem = (int)((sk2_warp_bytes(entries) + 127) & ~(size_t)127);
    plan->smem_bytes = (size_t)args.warp_smem * WARPS + 2 * WARPS * 4 + 16 + 8 + 8;
    int occ = 0;
    cudaError_t e;
    const bool mw = args.m > 16;
    if (args.canonical) e = mw ? sk2_setup<true, true>(plan->smem_bytes, &occ) : sk2_setup<true, false>(plan->smem_bytes, &occ);
    else e = mw ? sk2_setup<false, true>(plan->smem_bytes, &occ) : sk2_setup<false, false>(plan->smem_bytes, &occ);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
    long long grid = (long long)sm_count * occ;
    const long long units = (args.num_tiles + WARPS - 1) / WARPS;
    if (grid > units) grid = units;
    plan->grid = (int)(grid < 1 ? 1 : grid);
    return cudaSuccess;
}

cudaError_t sk2_launch(const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream)
{
    const bool mw = args.m > 16;
    if (args.canonical) {
        if (mw) sk2_kernel<true, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
        else sk2_kernel<true, false><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    } else {
        if (mw) sk2_kernel<false, true><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
        else sk2_kernel<false, false><<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(args);
    }
    return cudaGetLastError();
}

}  // namespace blg
