// sk4_device.cuh -- fused minimizer / super-k-mer extraction for sm_100a, fourth generation (device code; host side sk4_kernel.cu).
//
// Same results as sk3 / sk2 (SURVEY.md 8(a) decisions D1-D8; reference intent include/minimizer_view.hpp:184-266,
// include/super_kmer_view.hpp:121-135). What changed is again the instruction count. The ncu profile of sk3
// (profiles/r01_s27_sk3_kernel_instmix.txt) shows 160 executed instructions per base with the ALU pipe (LOP3 / SHF /
// ISETP / SEL / IADD3, 16 lanes per clock per scheduler) as the limiter, and only a third of them are the hash. Here
//
//   * the window minimum runs on ONE 32-bit key per m-mer: the top 26 bits of the hash above a 6-bit index (the m-mer's
//     position inside the lane's span). A running minimum is one VIMNMX; a tie in the hash part -- the only case where
//     26 bits do not decide what 64 bits decide -- is caught by one VIADDMNMX per compare (min over the key differences,
//     taken in the direction in which a tie is a small positive number). A tile that saw a tie (true duplicates in
//     low-complexity sequence, or a 2^-26 collision) is redone by exact_tile4() with full 64-bit compares, so the output
//     is bit-exact always; sk3 carried (hash hi, hash lo, index) through 2 ISETP + 3 SEL per compare;
//   * a lane owns NB consecutive blocks of B = w - 1 m-mer starts (its SPAN, L = NB B <= 32 starts), so a window never
//     reaches further than the next lane's first block: the index alone says where the minimum is (an index below the
//     window's own offset can only belong to the next lane), the prefix minima of all blocks but the first stay in
//     registers, and the three margin blocks of a tile (1 left, 2 right) are spread over 32 NB blocks instead of 32;
//   * everything per tile (tickets, ordered-offset scan, prefetch, warp scans) is paid once per 32 L starts.
//
// A warp owns a tile end to end and never meets another warp at a barrier, as in sk3: tiles come from a ticket counter,
// the ordered output offset from the two-level chained scan, records are written KB-1 tiles after they were computed.
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace blg {

namespace {

constexpr int NT4 = 32;
constexpr int WARPS4 = SK4_THREADS / 32;
constexpr int PADF4 = 3;     // zero words in front of the big-endian stream

template <int B, int NB, int Q = 1, int NP = 1>
struct Geo4 {
    // Q: a window is suffix-of-own-block + (Q-1 whole blocks) + prefix-of-block-(b+Q): w = Q B + 1 (Q = 1: k-m = 12, 10; Q = 2: k-m = 32)
    static constexpr int B_ = B;
    static constexpr int L = NB * B;                    // starts per lane (its span), <= 32
    static constexpr int S = NT4 * L;                   // m-mer starts hashed per tile
    static constexpr int OWN = NT4 * NB - 1 - 2 * Q;    // blocks 1 .. 32 NB - 1 - 2Q own record starts: a record reaches Q blocks
                                                        // further, and those windows Q blocks beyond that
    static constexpr int R = OWN * B;                   // new windows per tile
    static constexpr int NWORDS = S / 16 + 3;           // staged 16-base words: 15 (alignment) + S + 32 (m - 1) bases
    static constexpr int NROW = (NWORDS + 8) / 2 + 2;   // u32 words of the packed bad-byte / sequence-start bit rows
    static constexpr int BQ = (B + 3) / 4;              // uint4 per block of prefix keys
    // per-warp shared memory, byte offsets
    static constexpr int OFF_BAD = 0;                               // u32 [NROW]   bit x: base a0 + x is not ACGTU
    static constexpr int OFF_SEQ = OFF_BAD + NROW * 4;              // u32 [NROW]   bit x: a sequence starts at a0 + x
    static constexpr int OFF_FLAG = (OFF_SEQ + NROW * 4 + 7) / 8 * 8;  // int [2], then the warp's mbarrier (8 bytes)
    static constexpr int PBLK = Q == 1 ? 1 : NB;                    // blocks per lane whose prefix keys another lane reads
    static constexpr int OFF_P = (OFF_FLAG + 16 + 15) / 16 * 16;    // uint4 [PBLK][BQ][32] (+1)  prefix keys, [block][quad][lane]
    static constexpr int OFF_IDX2 = OFF_P + NP * (PBLK * BQ * 32 + 1) * 16;   // NP = 2 (syncmers, both strands): a second row of prefix keys
    static constexpr int OFF_BUF = OFF_IDX2 + (NP - 1) * 32 * 32;               // and of argmin bytes, u8 [32][2][16]
    // per buffer set (tile i is written out while tiles i+1, i+2 are computed)
    static constexpr int BUF_CW = 0;                                // u32 [NWORDS + 3]  complemented codes, little-endian
    static constexpr int BUF_EW = BUF_CW + (NWORDS + 3) * 4;        // u32 [PADF4 + NWORDS + 3]  codes, big-endian words
    static constexpr int BUF_IDX = (BUF_EW + (PADF4 + NWORDS + 3) * 4 + 15) / 16 * 16;   // u8 [32][2][16]  argmin byte of every window, [lane][block][window]
    static constexpr int BUF_EBW = BUF_IDX + 32 * 32;               // u32 [33 + 3]  record-end bits of each lane's span
    static constexpr int BUF_REC = BUF_EBW + 36 * 4;                // u16 [R + 8]  record starts (lane << 5 | j)
    static constexpr int BUF_BYTES = (BUF_REC + (R + 8) * 2 + 15) / 16 * 16;
#ifdef SK4_KB
    static constexpr int KB = SK4_KB;
#else
    static constexpr int KB = Q == 1 ? 3 : 2;                       // Q = 2: three sets would leave 12 warps per SM
#endif
    static constexpr int WARP_BYTES = (OFF_BUF + KB * BUF_BYTES + 127) / 128 * 128;
    // the next tile's ASCII bytes (16 NWORDS) land in the REC area of the buffer set the tile will be staged into; the
    // area is free from the end of that set's emit until the tile's own record slots are written, after staging has consumed the bytes
    static constexpr int RAW_BYTES = 16 * NWORDS;
    static_assert(BUF_REC % 16 == 0 && (R + 8) * 2 >= RAW_BYTES, "the raw ASCII of a tile must fit the record-slot area");
};

__device__ __forceinline__ u32 pairrev32_4(u32 x)
{
    x = __brev(x);
    return ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
}

// (h & mask) | idx in one LOP3: the mask is a kernel argument so that it stays a register / constant-bank operand
__device__ __forceinline__ u32 make_key(u32 h, u32 mask, u32 idx)
{
    u32 k;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(k) : "r"(h), "r"(mask), "r"(idx));
    return k;
}

// low NBITS bits: OR over d in [0, span) of bit (bitpos + i + d + shift) of the packed bit row, for i = 0 .. NBITS-1
// (NBITS + span + shift <= 96)
template <int NBITS>
__device__ __forceinline__ u32 or_span_bits4(const u32* row, int bitpos, int span, int shift)
{
    const int wq = bitpos >> 5, sh = bitpos & 31;
    const u32 w0 = row[wq], w1 = row[wq + 1], w2 = row[wq + 2], w3 = row[wq + 3];
    u32 x0 = __funnelshift_r(w0, w1, sh), x1 = __funnelshift_r(w1, w2, sh), x2 = __funnelshift_r(w2, w3, sh);
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
    return NBITS >= 32 ? x0 : (x0 & ((1u << (NBITS & 31)) - 1u));
}

struct TileConst4 {
    u32 maskhi, masklo, keymask, s8;
    int w, m, k;
};

template <int N>
__device__ __forceinline__ u32 word_or_zero(const u32 (&a)[N], int i) { return (i >= 0 && i < N) ? a[i] : 0u; }

// canonical-or-forward m-mer value at local base index xj, cut out of the staged streams with run-time shifts
template <bool CANON, bool MW>
__device__ __forceinline__ u64 mmer_from_smem(const u32* CW, const u32* EW, int xj, const TileConst4& c)
{
    u64 val;
    {
        const int e = 2 * xj + 2 * c.m + 32 * PADF4 - 64;     // the 64 bits that end with the m-mer's last base
        const int wj = e >> 5, fs = e & 31;
        const u32 e0 = EW[wj], e1 = EW[wj + 1], e2 = EW[wj + 2];
        u32 fhi = __funnelshift_l(e1, e0, fs), flo = __funnelshift_l(e2, e1, fs);
        if (MW) fhi &= c.maskhi; else { fhi = 0; flo &= c.masklo; }
        val = ((u64)fhi << 32) | flo;
    }
    if (CANON) {
        const int wi = xj >> 4, sh = (xj & 15) * 2;
        const u32 c0 = CW[wi], c1 = CW[wi + 1], c2 = CW[wi + 2];
        u32 rlo = __funnelshift_r(c0, c1, sh), rhi = __funnelshift_r(c1, c2, sh);
        if (MW) rhi &= c.maskhi; else { rhi = 0; rlo &= c.masklo; }
        const u64 rc = ((u64)rhi << 32) | rlo;
        val = rc < val ? rc : val;
    }
    return val;
}

// Exact redo of the windows a tie could have changed (full 64-bit hash compares, leftmost on ties: the strict '<' of
// reference include/minimizer_view.hpp:283,374). `flagged` = lanes whose compares saw equal hash parts. A tie seen by
// lane f can only concern windows [L f - B, L f + L) (its own span, and the previous lane's last block, which reads lane
// f's block-0 prefix). The warp redoes them cooperatively in UNITS of NU = 33 - w windows: lane j hashes m-mer start
// s + j, window s + j scans hashes j .. j + w - 1 through shuffles. A unit costs about 200 instructions, so the usual
// case -- one equal pair of m-mers in a tile, e.g. reverse-complement neighbours at small m -- adds a fraction of a
// tile's time, and a tile full of ties (low-complexity sequence) about four tile times. Overwrites the argmin bytes
// of the redone windows in IDXB; the caller rebuilds its change bits from IDXB afterwards (change_bits_from_idx4).
template <int B, int NB, int Q, bool CANON, bool MW>
__device__ __noinline__ void redo_windows4(const u32* CW, const u32* EW, unsigned char* IDXB, TileConst4 c, int o, int t, u32 flagged)
{
    constexpr int L = NB * B, W = Q * B + 1, S = 32 * L;
    constexpr int H = W <= 17 ? 1 : 2;                 // m-mer starts hashed per lane and unit
    constexpr int NU = 32 * H - W + 1, UNITS = (S + NU - 1) / NU;
    static_assert(UNITS <= 64, "unit mask is 64 bits");
    // units to redo: those that overlap [L f - Q B, L f + L) of a flagged lane f (the lanes whose blocks read f's prefix keys)
    u64 todo = 0;
    for (u32 fm = flagged; fm; fm &= fm - 1) {
        const int f = __ffs((int)fm) - 1;
        int lo = L * f - Q * B, hi = L * f + L;
        lo = lo < 0 ? 0 : lo;
        const int u0 = lo / NU, u1 = (hi - 1) / NU;
        for (int u = u0; u <= u1 && u < UNITS; ++u) todo |= 1ULL << u;
    }
    for (; todo; todo &= todo - 1) {
        const int s = (__ffsll((long long)todo) - 1) * NU;
        u64 h[H];
#pragma unroll
        for (int q = 0; q < H; ++q) h[q] = murmur3_key8_s32(mmer_from_smem<CANON, MW>(CW, EW, o + s + 32 * q + t, c), c.s8);
        u64 best = h[0];
        int bd = 0;
#pragma unroll 4
        for (int d = 1; d < W; ++d) {
            const int src = t + d;                     // hash of start s + t + d: lane src & 31 of h[src >> 5]
            u64 hd = __shfl_sync(0xffffffffu, h[0], src & 31);
            if (H == 2) { const u64 h1 = __shfl_sync(0xffffffffu, h[H - 1], src & 31); if (src >= 32) hd = h1; }
            if (hd < best) { best = hd; bd = d; }
        }
        const int ws = s + t;                          // this lane's window
        if (t < NU && ws < S) {
            const int gb = ws / B, i = ws - gb * B;    // block of the tile (= lane * NB + block in lane), window in block
            const int pb = i + bd;                      // minimizer start relative to the block's first start: up to Q blocks on
            const int pv = pb / B;
            IDXB[gb * 16 + i] = (unsigned char)((pv << 4) | (pb - pv * B));
        }
    }
    __syncwarp();
}

// The lane's change bits (bit j: windows j and j+1 of the span have different minimizer positions) from the argmin bytes
// in shared memory; only used after redo_windows4.
template <int B, int NB>
__device__ __forceinline__ u32 change_bits_from_idx4(const unsigned char* IDXB, int t)
{
    constexpr int L = NB * B;
    const uint4* IW = reinterpret_cast<const uint4*>(IDXB + t * 32);
    u32 wv[NB * 4];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) { const uint4 e = IW[nb]; wv[4 * nb] = e.x; wv[4 * nb + 1] = e.y; wv[4 * nb + 2] = e.z; wv[4 * nb + 3] = e.w; }
    const int n0 = IDXB[(t + 1) * 32];                  // lane 31 reads the first byte behind the array: margin, any value
    u32 dif = 0;
    int pcur = 0;
#pragma unroll
    for (int j = 0; j <= L; ++j) {
        // position of window j's minimizer relative to the lane's first start (window L = the next lane's window 0)
        const int nb = j / B, i = j - nb * B;
        const int byt = j < L ? (int)((wv[4 * nb + (i >> 2)] >> (8 * (i & 3))) & 63u) : (n0 & 63);
        const int pos = nb * B + (byt & 15) + ((byt >> 4) & 3) * B;
        if (j > 0) dif |= (u32)(pos != pcur) << (j - 1);
        pcur = pos;
    }
    return dif;
}

// The lane's selected windows (bit j of `start`: window j of its span) -> their slots in REC, in tile order: a window is named by
// its slot in the argmin array, lane << 5 | block-in-lane << 4 | window-in-block. Returns the tile's count.
template <int B>
__device__ __forceinline__ u32 record_slots4(unsigned short* REC, u32 start, int t)
{
    static_assert(B <= 16, "slot numbering assumes two blocks of at most 16 windows per lane");
    constexpr u32 MB = B >= 32 ? 0xFFFFFFFFu : ((1u << B) - 1u);
    const u32 cnt = __popc(start);
    u32 incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 v = __shfl_up_sync(0xffffffffu, incl, d);
        if (t >= d) incl += v;
    }
    unsigned short* rp = REC + (incl - cnt);
    u32 sm = B == 16 ? start : ((start & MB) | ((start >> B) << 16));
    const u32 tag = (u32)t << 5;
    while (sm) {
        const int i = __ffs((int)sm) - 1;
        sm &= sm - 1;
        *rp++ = (unsigned short)(tag | (u32)i);
    }
    __syncwarp();
    return __shfl_sync(0xffffffffu, incl, 31);
}

// Stage one tile into buffer set b: ASCII -> the two packed streams + bad-byte bits + sequence-start bits. Returns the valid mask
// of the lane's L windows (bit j: the k-mer at the lane's start j has k good bases of one sequence); *o_out = offset of local start 0
// from the 16-base boundary the staged range begins at.
template <int B, int NB, int Q, int NP>
__device__ __forceinline__ u32 stage_tile4(const SkArgs& a, unsigned char* wsm, const TileConst4& c, int b, long long tile, int t,
                                           int have_pre, int* o_out)
{
    using G = Geo4<B, NB, Q, NP>;
    constexpr int L = G::L;
    constexpr u32 ML = L >= 32 ? 0xFFFFFFFFu : ((1u << (L & 31)) - 1u);
    const long long g0 = a.org + tile * (long long)G::R - B;     // global position of local start 0 (block 0 = left margin)
    const long long a0 = g0 & ~15LL;                              // staged range starts on a 16-base boundary
    const int o = (int)(g0 - a0);
    *o_out = o;
    unsigned char* buf = wsm + G::OFF_BUF + b * G::BUF_BYTES;
    u32* BAD = reinterpret_cast<u32*>(wsm + G::OFF_BAD);
    u32* SEQ = reinterpret_cast<u32*>(wsm + G::OFF_SEQ);
    int* FLAG = reinterpret_cast<int*>(wsm + G::OFF_FLAG);
    u32* CW = reinterpret_cast<u32*>(buf + G::BUF_CW);
    u32* EW = reinterpret_cast<u32*>(buf + G::BUF_EW);
    u32* EBW = reinterpret_cast<u32*>(buf + G::BUF_EBW);
    // ---- stage: ASCII -> packed streams + bad-byte bits -----------------------------------------------------------
    int anybad = 0;
    // have_pre: bit 0 = the tile's bytes were requested by a bulk async copy (tile_loop4), bit 1 = the parity of that copy's mbarrier
    // phase. They sit in the record-slot area of this buffer set until the record slots are written at the end of the tile.
    const unsigned char* raw = buf + G::BUF_REC;
    if (have_pre) mbar_wait(reinterpret_cast<unsigned long long*>(wsm + G::OFF_FLAG + 8), (u32)(have_pre >> 1) & 1u);
    {
        unsigned short* BAD16 = reinterpret_cast<unsigned short*>(BAD);
#pragma unroll
        for (int wi = t; wi < 2 * G::NROW; wi += NT4) {
            u32 codes = 0, bad = 0xFFFFu;
            if (wi < G::NWORDS) {
                const long long P16 = a0 + 16LL * wi;
                u32 w0, w1, w2, w3;
                if (have_pre) {
                    const uint4 v = *reinterpret_cast<const uint4*>(raw + 16 * wi);
                    w0 = v.x; w1 = v.y; w2 = v.z; w3 = v.w;
                } else if (a.vec_ok && P16 >= a.buf_lo && P16 + 16 <= a.buf_hi) {
                    const uint4 v = __ldg(reinterpret_cast<const uint4*>(a.bases + (P16 - a.buf_lo)));
                    w0 = v.x; w1 = v.y; w2 = v.z; w3 = v.w;
                } else {
                    w0 = w1 = w2 = w3 = 0;                 // tile edges / unaligned buffers: byte loads, rolled
#pragma unroll 1
                    for (int q = 0; q < 16; ++q) {
                        const long long pos = P16 + q;
                        const u32 byte = (pos >= a.buf_lo && pos < a.buf_hi) ? (u32)a.bases[pos - a.buf_lo] : 0u;
                        const u32 sh = byte << (8 * (q & 3));
                        if (q < 4) w0 |= sh; else if (q < 8) w1 |= sh; else if (q < 12) w2 |= sh; else w3 |= sh;
                    }
                }
                // fast path: codes + one exact "any byte invalid" word; the per-base bits only where a word has a bad byte
                if ((blg_bad4(w0) | blg_bad4(w1) | blg_bad4(w2) | blg_bad4(w3)) == 0u) {
                    codes = blg_codes4(w0) | (blg_codes4(w1) << 8) | (blg_codes4(w2) << 16) | (blg_codes4(w3) << 24);
                    bad = 0;
                } else {
                    blg_pack16_bits(w0, w1, w2, w3, &codes, &bad);
                    anybad = 1;
                }
            }
            if (wi < G::NWORDS + 3) {
                CW[wi] = ~codes;
                EW[PADF4 + wi] = pairrev32_4(codes);
            }
            BAD16[wi] = (unsigned short)bad;
        }
        if (t < PADF4) EW[t] = 0;
        for (int x = t; x < G::NROW; x += NT4) SEQ[x] = 0;
        if (t == 0) { FLAG[0] = 0; EBW[32] = 0; }
    }
    __syncwarp();
    const int dirty = __any_sync(0xffffffffu, anybad);
    if (a.seq_offsets != nullptr) {
        // interior sequence starts inside the staged range
        const long long lo_pos = a0 < 0 ? 0 : a0, end_pos = a0 + 16LL * G::NWORDS;
        const long long lo = first_offset_at_or_after(a.seq_offsets, a.n_seq, lo_pos, a.buf_hi, t);
        int found = 0;
        for (long long q = lo + t; q < a.n_seq; q += 32) {
            const long long off = (long long)a.seq_offsets[q];
            if (off >= end_pos) break;
            const int rel = (int)(off - a0);
            atomicOr(&SEQ[rel >> 5], 1u << (rel & 31));
            found = 1;
        }
        if (found) FLAG[0] = 1;
        __syncwarp();
    }
    const int seqdirty = (a.seq_offsets != nullptr) && FLAG[0];

    // ---- validity of the lane's windows (before the minima: a lane without a valid window cannot force the exact path) --
    const int x0 = o + L * t;                                  // base index of this lane's first start, from a0
    u32 V = ML;
    if (dirty | seqdirty) {
        u32 inv = 0;
        if (dirty) inv |= or_span_bits4<L>(BAD, x0, c.k, 0);
        if (seqdirty && c.k > 1) inv |= or_span_bits4<L>(SEQ, x0, c.k - 1, 1);
        V = ~inv & ML;
    }
    return V;
}

// Stage + hash + window minima + record starts of one tile into buffer set b. Returns the tile's record count.
template <int B, int NB, int Q, int LW, bool CANON, bool MW>
__device__ __forceinline__ u32 compute_tile4(const SkArgs& a, unsigned char* wsm, const TileConst4& c, int b, long long tile, int t,
                                             int have_pre)
{
    using G = Geo4<B, NB, Q>;
    constexpr int L = G::L;
    constexpr u32 ML = L >= 32 ? 0xFFFFFFFFu : ((1u << (L & 31)) - 1u);
    const long long g0 = a.org + tile * (long long)G::R - B;
    int o;
    const u32 V = stage_tile4<B, NB, Q, 1>(a, wsm, c, b, tile, t, have_pre, &o);
    unsigned char* buf = wsm + G::OFF_BUF + b * G::BUF_BYTES;
    uint4* P = reinterpret_cast<uint4*>(wsm + G::OFF_P);
    u32* CW = reinterpret_cast<u32*>(buf + G::BUF_CW);
    u32* EW = reinterpret_cast<u32*>(buf + G::BUF_EW);
    unsigned char* IDXB = buf + G::BUF_IDX;
    u32* EBW = reinterpret_cast<u32*>(buf + G::BUF_EBW);
    unsigned short* REC = reinterpret_cast<unsigned short*>(buf + G::BUF_REC);
    const int x0 = o + L * t;                                  // base index of this lane's first start, from the staged range's start

    // ---- the lane's blocks, left to right (a ROLLED loop: the unrolled body of one block is what the instruction cache holds).
    // Iteration nb hashes block nb -- one 32-bit key per m-mer start (hash bits above the index INSIDE the block), running
    // prefix minimum pk[] -- and, for Q = 1, finishes the windows of block nb-1: suffix minimum of its own keys combined with
    // pk[] (a window is suffix-of-own-block + [Q-1 whole blocks] + prefix-of-block-(b+Q)). Keys that come from a later block
    // enter with +16 per block, so the low byte of a window's minimum says where it is: bits 0-3 index in its block, bits 4-5
    // how many blocks further on.
    u32 tacc = 0xFFFFFFFFu;                    // min over all compares of (right key - left key): below 32 <=> a hash-part tie
    u32 dif = 0;                               // bit j: windows j and j+1 of the span have different minimizer positions
    u32 ks_prev[B], ks_prev2[B];               // keys of the last and (Q = 2) the last but one hashed block
    u32 full_prev = 0;                         // minimum key of the last hashed block (Q = 2)
    u32 mv_first = 0, mv_last = 0;
    // windows of one block: own keys kk[] (suffix side), for Q = 2 the minimum key of the block after it, and the prefix keys
    // np[] of the block Q further on; returns the block's change bits (bit i: windows i, i+1; bit B-1 is set by the caller)
    // and stores the argmin bytes
    auto finish_block = [&](const u32 (&kk)[B], u32 fullkey, const u32 (&np)[G::BQ * 4], int nbp) -> u32 {
        u32 mvb[B];
        u32 sm = 0;
#pragma unroll
        for (int i = B - 1; i >= 0; --i) {
            if (i == B - 1) {
                if (Q == 1) sm = kk[i];
                else {
                    sm = fullkey + 16u; tacc = min(tacc, sm - kk[i]);
                    sm = min(kk[i], sm);
                }
            } else {
                tacc = min(tacc, sm - kk[i]);
                sm = min(kk[i], sm);
            }
            tacc = min(tacc, sm - np[i]);              // equal hash parts: what sm holds lies left of what np holds: a small difference >= 0
            mvb[i] = min(np[i] + 16u * Q, sm);
        }
        u32 wv[4] = {0, 0, 0, 0};
#pragma unroll
        for (int q = 0; q < (B + 3) / 4; ++q) {
            const u32 m0 = mvb[4 * q], m1 = 4 * q + 1 < B ? mvb[4 * q + 1] : 0u, m2 = 4 * q + 2 < B ? mvb[4 * q + 2] : 0u,
                      m3 = 4 * q + 3 < B ? mvb[4 * q + 3] : 0u;
            wv[q] = __byte_perm(__byte_perm(m0, m1, 0x0040), __byte_perm(m2, m3, 0x0040), 0x5410);
        }
        *reinterpret_cast<uint4*>(IDXB + t * 32 + nbp * 16) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
        // change bits inside the block, four windows per step on the packed bytes: two windows of one block have the same
        // minimizer iff their (blocks-further-on, index) fields are equal. Byte i against byte i+1: x = low six bits of the xor;
        // (x + 0x7F) & 0x80 flags a non-zero byte; the multiply gathers the four flags into the top nibble.
        u32 d = 0;
#pragma unroll
        for (int q = 0; q < (B + 3) / 4; ++q) {
            const u32 nxt = __funnelshift_r(wv[q], q + 1 < 4 ? wv[q + 1] : 0u, 8);
            const u32 x = (wv[q] ^ nxt) & 0x3F3F3F3Fu;
            const u32 nz = (x + 0x7F7F7F7Fu) & 0x80808080u;
            d |= ((nz * 0x00204081u) >> 28) << (4 * q);
        }
        d &= (1u << (B - 1)) - 1u;                     // bit B-1 (this block's last window against the next block's first) is the caller's
        // the previous block's last window against this block's first: the same position reads +16 from there
        if (nbp > 0) dif |= (u32)(mvb[0] + 16u != mv_last) << (nbp * B - 1);
        else mv_first = mvb[0];
        mv_last = mvb[B - 1];
        return d;
    };
#pragma unroll 1
    for (int nb = 0; nb < NB; ++nb) {
        // block-local streams: LW words from the block's first base (complemented little-endian side), and the LW words
        // that END with the last base of the block's last m-mer (big-endian side)
        const int xb = x0 + nb * B;
        u32 Rl[LW], Fl[LW];
        {
            const int wi0 = xb >> 4, sh = (xb & 15) * 2;
            u32 cw[LW + 1];
#pragma unroll
            for (int j = 0; j <= LW; ++j) cw[j] = CW[wi0 + j];
#pragma unroll
            for (int j = 0; j < LW; ++j) Rl[j] = __funnelshift_r(cw[j], cw[j + 1], sh);   // Rl[0] = least significant
            const int bp = 2 * (xb + B - 1 + c.m) - 32 * LW + 32 * PADF4;
            const int wj = bp >> 5, fs = bp & 31;
            u32 ew[LW + 1];
#pragma unroll
            for (int j = 0; j <= LW; ++j) ew[j] = EW[wj + j];
#pragma unroll
            for (int j = 0; j < LW; ++j) Fl[j] = __funnelshift_l(ew[j + 1], ew[j], fs);   // Fl[0] = most significant
        }
        u32 ks[B];
        u32 pk[G::BQ * 4];
        u32 pm = 0;
#pragma unroll
        for (int i = 0; i < B; ++i) {
            u32 flo, fhi = 0;
            {
                const int cI = 2 * (B - 1 - i), wq = cI >> 5, sh = cI & 31;
                flo = __funnelshift_r(word_or_zero(Fl, LW - 1 - wq), word_or_zero(Fl, LW - 2 - wq), sh);
                if (MW) fhi = __funnelshift_r(word_or_zero(Fl, LW - 2 - wq), word_or_zero(Fl, LW - 3 - wq), sh) & c.maskhi;
                else flo &= c.masklo;
            }
            u64 val = ((u64)fhi << 32) | flo;
            if (CANON) {
                const int bo = 2 * i, wq = bo >> 5, sh = bo & 31;
                u32 rlo = __funnelshift_r(word_or_zero(Rl, wq), word_or_zero(Rl, wq + 1), sh), rhi = 0;
                if (MW) rhi = __funnelshift_r(word_or_zero(Rl, wq + 1), word_or_zero(Rl, wq + 2), sh) & c.maskhi;
                else rlo &= c.masklo;
                const u64 rc = ((u64)rhi << 32) | rlo;
                val = rc < val ? rc : val;
            }
            const u32 hh = (u32)(murmur3_key8_s32(val, c.s8) >> 32);
            ks[i] = make_key(hh, c.keymask, (u32)i);
            if (i == 0) pm = ks[i];
            else {
                tacc = min(tacc, ks[i] - pm);
                pm = min(pm, ks[i]);
            }
            pk[i] = pm;
        }
#pragma unroll
        for (int i = B; i < G::BQ * 4; ++i) pk[i] = 0;
        if (nb < G::PBLK) {
#pragma unroll
            for (int q = 0; q < G::BQ; ++q) P[(nb * G::BQ + q) * 32 + t] = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
        }
        if (Q == 1 && nb >= 1) dif |= finish_block(ks_prev, 0u, pk, nb - 1) << ((nb - 1) * B);
#pragma unroll
        for (int i = 0; i < B; ++i) { ks_prev2[i] = ks_prev[i]; ks_prev[i] = ks[i]; }
        full_prev = pm;
    }
    __syncwarp();
    if (Q == 1) {
        // the last block: its next block is the next lane's block 0
        u32 np[G::BQ * 4];
#pragma unroll
        for (int q = 0; q < G::BQ; ++q) {
            const uint4 e = P[q * 32 + t + 1];
            np[4 * q] = e.x; np[4 * q + 1] = e.y; np[4 * q + 2] = e.z; np[4 * q + 3] = e.w;
        }
        dif |= finish_block(ks_prev, 0u, np, NB - 1) << ((NB - 1) * B);
    } else {
        // Q = 2, two blocks per lane: block 0 = own suffix + own block 1 + the next lane's block-0 prefix; block 1 = own suffix +
        // the next lane's block 0 + its block-1 prefix
        static_assert(Q == 1 || NB == 2, "Q = 2 is written for two blocks per lane");
        u32 np[G::BQ * 4];
#pragma unroll
        for (int q = 0; q < G::BQ; ++q) {
            const uint4 e = P[q * 32 + t + 1];
            np[4 * q] = e.x; np[4 * q + 1] = e.y; np[4 * q + 2] = e.z; np[4 * q + 3] = e.w;
        }
        const u32 next_full0 = np[B - 1];
        dif |= finish_block(ks_prev2, full_prev, np, 0);
#pragma unroll
        for (int q = 0; q < G::BQ; ++q) {
            const uint4 e = P[(G::BQ + q) * 32 + t + 1];
            np[4 * q] = e.x; np[4 * q + 1] = e.y; np[4 * q + 2] = e.z; np[4 * q + 3] = e.w;
        }
        dif |= finish_block(ks_prev, next_full0, np, 1) << B;
    }
    {
        const u32 nv0 = __shfl_down_sync(0xffffffffu, mv_first, 1);
        dif |= (u32)(nv0 + 16u != mv_last) << (L - 1);
    }
#ifndef SK4_NO_EXACT     // timing experiments only: without the redo a tile with ties may be wrong
    {
        // Q = 2: lane 31 holds the last two margin blocks; no owned record reaches its windows, and what it reads as "next lane"
        // is other lanes' keys (a seed and a prefix entry can be the same key there): it never asks for a redo
        const u32 flagged = __ballot_sync(0xffffffffu, tacc < 32u && V != 0u && (Q == 1 || t < 31));
        if (flagged) {
            __syncwarp();
            redo_windows4<B, NB, Q, CANON, MW>(CW, EW, IDXB, c, o, t, flagged);
            dif = change_bits_from_idx4<B, NB>(IDXB, t);
            if (t == 0 && a.exact_tiles != nullptr) atomicAdd(a.exact_tiles, 1ULL);
        }
    }
#endif

    // ---- record ends / starts -----------------------------------------------------------------------------------------
    const u32 Vn = (V >> 1) | ((__shfl_down_sync(0xffffffffu, V, 1) & 1u) << (L - 1));
    const u32 E = V & (~Vn | dif) & ML;        // the record ends with this window
    const u32 X = (E | ~V) & ML;               // the next window cannot continue a record
    EBW[t] = E;
    const u32 xprev = (__shfl_up_sync(0xffffffffu, X, 1) >> (L - 1)) & 1u;
    u32 start = V & ((X << 1) | xprev) & ML;
    {
        // owned windows: local [B, B + R), clipped to the call's [win_lo, win_hi)
        long long hi_g = a.win_hi - g0;
        const int hi_s = hi_g > (long long)(B + G::R) ? B + G::R : (int)hi_g;
        int lo_j = B - L * t, hi_j = hi_s - L * t;
        lo_j = lo_j < 0 ? 0 : (lo_j > L ? L : lo_j);
        hi_j = hi_j < 0 ? 0 : (hi_j > L ? L : hi_j);
        const u32 lom = lo_j >= 32 ? 0u : (0xFFFFFFFFu << lo_j);
        const u32 him = hi_j >= 32 ? 0xFFFFFFFFu : ((1u << hi_j) - 1u);
        start &= lom & him;
    }
    return record_slots4<B>(REC, start, t);
}

// Record-parallel output of a finished tile from buffer set b: `n_rec` records, the first one at global index `base`.
template <int B, int NB, int Q, bool CANON, bool MW>
__device__ __forceinline__ void emit_tile4(const SkArgs& a, const unsigned char* wsm, const TileConst4& c, int b, long long tile, int t,
                                           u32 n_rec, u64 base)
{
    using G = Geo4<B, NB, Q>;
    constexpr int L = G::L;
    const long long g0 = a.org + tile * (long long)G::R - B;
    const int o = (int)(g0 - (g0 & ~15LL));
    const unsigned char* buf = wsm + G::OFF_BUF + b * G::BUF_BYTES;
    const u32* CW = reinterpret_cast<const u32*>(buf + G::BUF_CW);
    const u32* EW = reinterpret_cast<const u32*>(buf + G::BUF_EW);
    const unsigned char* IDXB = buf + G::BUF_IDX;
    const u32* EBW = reinterpret_cast<const u32*>(buf + G::BUF_EBW);
    const unsigned short* REC = reinterpret_cast<const unsigned short*>(buf + G::BUF_REC);
    for (u32 r = t; r < n_rec; r += NT4) {
        const u32 rec = REC[r];
        const int t0 = (int)(rec >> 5), nb0 = (int)((rec >> 4) & 1u), i0 = (int)(rec & 15u);
        const int j0 = nb0 * B + i0;                            // the window's offset in its lane's span
        const int byt = IDXB[rec];                              // bits 0-3: index in its block, bits 4-5: how many blocks further on
        const int rel = (byt & 15) + ((byt >> 4) & 3) * B - i0; // mm_pos
        const int j = L * t0 + j0 + rel;                        // local m-mer start
        // size: distance to the first record-end bit at or after the start, at most w
        int size;
        {
            const u32 e0 = EBW[t0], e1 = EBW[t0 + 1];
            const u32 bits = L == 32 ? __funnelshift_r(e0, e1, j0) : ((e0 >> j0) | (e1 << ((L - j0) & 31)));
            size = bits ? __ffs((int)bits) : c.w;
            size = size < c.w ? size : c.w;
        }
        const u64 val = mmer_from_smem<CANON, MW>(CW, EW, o + j, c);
        const u64 g = base + r;
        if (g < a.capacity) {
            a.value[g] = val;
            store_mm_start(a, g, g0 + j);
            if (a.size != nullptr) { a.size[g] = (u8)size; a.mm_pos[g] = (u8)rel; }
            if (a.hash != nullptr) a.hash[g] = murmur3_key8_s32(val, c.s8);
        }
    }
    __syncwarp();   // this buffer set is restaged by a later compute_tile
}

// ---- ordered output offset without any CTA barrier: the two-level chained scan of sk3 (tiles -> groups of 32) ------------
// No scratch word is zeroed between launches: every word carries the launch's EPOCH (a.epoch, 22 bits, bumped by the host
// per launch) and a word of another epoch reads as "not written yet"; tickets are drawn from a counter that only grows
// (a.ticket_base = draws of all earlier launches: each launch draws num_tiles + number-of-warps tickets); a group's
// arrival accumulator is reset by the warp that completes the group. The host zeroes the scratch when it allocates it
// and when the epoch wraps.
//   tile word  : (1 << 31 | epoch) << 32 | records of the tile
//   group word : flag (1 = aggregate, 2 = inclusive prefix) << 62 | epoch << 40 | count (40 bits)
constexpr int GS4 = 32;
constexpr u64 ACC_ONE4 = 1ULL << 40, ACC_MASK4 = ACC_ONE4 - 1;
constexpr u64 CNT_MASK4 = (1ULL << 40) - 1;

__device__ __forceinline__ void spin_wait4(u32& spins)
{
    __nanosleep(64);
#ifdef BLG_DEBUG_SPIN
    if (++spins > (1u << 28)) __trap();   // debug builds only: a protocol bug must not hang the GPU
#else
    (void)spins;
#endif
}

__device__ __forceinline__ u32 tile_tag4(const SkArgs& a) { return 0x80000000u | a.epoch; }
__device__ __forceinline__ u32 grp_tag4(const SkArgs& a, u32 flag) { return (flag << 22) | a.epoch; }   // compares with word >> 40

// decoupled look-back over the group words; returns the exclusive prefix of group g and publishes its inclusive prefix
__device__ __forceinline__ u64 lookback_groups4(const SkArgs& a, long long g, u64 aggregate, int lane)
{
    const unsigned full = 0xffffffffu;
    u64* status = a.grp_status;
    const u64 ep = (u64)a.epoch << 40;
    if (g == 0) {
        if (lane == 0) st_status(&status[0], (2ULL << 62) | ep | aggregate);
        return 0;
    }
    if (lane == 0) st_status(&status[g], (1ULL << 62) | ep | aggregate);
    const u32 t_agg = grp_tag4(a, 1), t_inc = grp_tag4(a, 2);
    u64 excl = 0;
    long long look = g - 1;
    for (;;) {
        const long long idx = look - lane;
        u64 st = idx >= 0 ? ld_status(&status[idx]) : ((2ULL << 62) | ep);
        u32 tag = (u32)(st >> 40);
        while (__any_sync(full, tag != t_agg && tag != t_inc)) {
            if (tag != t_agg && tag != t_inc) { st = ld_status(&status[idx]); tag = (u32)(st >> 40); }
        }
        const unsigned incl = __ballot_sync(full, tag == t_inc);
        const int first_incl = incl ? (__ffs(incl) - 1) : 32;
        u64 cnt = lane <= first_incl ? (st & CNT_MASK4) : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(full, cnt, o);
        excl += cnt;
        if (incl) break;
        look -= 32;
    }
    if (lane == 0) st_status(&status[g], (2ULL << 62) | ep | (excl + aggregate));
    return excl;
}

__device__ __forceinline__ const u64* tile_base_addr4(const SkArgs& a, long long tile, int t)
{
    const long long g = tile / GS4;
    const int idx = (int)(tile - g * GS4);
    if (t < idx) return a.tile_status + g * GS4 + t;
    if (t == 31 && g > 0) return a.grp_status + (g - 1);
    return nullptr;
}
// lanes below the tile's index in its group hold the earlier tiles' counts (small), lane 31 the inclusive prefix of the
// previous group (40 bits): one REDUX for the counts, one 64-bit broadcast
__device__ __forceinline__ u64 tile_base_finish4(const SkArgs& a, const u64* p, u64 v, int t)
{
    if (p != nullptr) {
        u32 spins = 0;
        if (t == 31) {
            const u32 want = grp_tag4(a, 2);
            while ((u32)(v >> 40) != want) { spin_wait4(spins); v = ld_status(p); }
            v &= CNT_MASK4;
        } else {
            const u32 want = tile_tag4(a);
            while ((u32)(v >> 32) != want) { spin_wait4(spins); v = ld_status(p); }
            v &= 0xFFFFFFFFULL;
        }
    }
    const u32 small = t == 31 ? 0u : (u32)v;
    const u32 s = __reduce_add_sync(0xffffffffu, small);
    const u64 g = __shfl_sync(0xffffffffu, v, 31);
    return g + s;
}

// The loop every warp runs on its own: compute tile i (`compute(b, tile, have_pre)` -> its record count), publish the count,
// write out tile i-(KB-1) from its buffer set (`emit(b, tile, n_rec, base)`). A ticket is taken only when nothing can block any
// more before that tile is published (a warp that waits for a prefix while it holds an uncomputed tile stalls that tile's group,
// the groups behind it, and so itself: convoys).
template <class G, class ComputeF, class EmitF>
__device__ __forceinline__ void tile_loop4(const SkArgs& a, unsigned char* wsm, int t, ComputeF compute, EmitF emit)
{
    const long long num_groups = (a.num_tiles + GS4 - 1) / GS4;
    unsigned long long* mbar = reinterpret_cast<unsigned long long*>(wsm + G::OFF_FLAG + 8);   // completion of the warp's bulk copies
    if (t == 0) mbar_init(mbar, 1);
    __syncwarp();
    u32 bulk_phase = 0;
    u32 tk = 0;
    if (t == 0) tk = atomicAdd(a.tile_counter, 1u) - a.ticket_base;
    long long tile = __shfl_sync(0xffffffffu, tk, 0);
    // computed, not yet written out: (t0, n0, b0) the newer, (t1, n1, b1) the older (KB == 3 only): tile, records, buffer set
    long long t0 = -1, t1 = -1;
    u32 n0 = 0, n1 = 0;
    int b0 = 0, b1 = 0;
    int have_pre = 0;
    int b = 0;                          // buffer set of the tile computed in this iteration
    for (;;) {
        const bool have = tile < a.num_tiles;
        // the tile to write out now: the older pending one once two are queued, or whatever is left when draining
        const bool from1 = G::KB == 3 && t1 >= 0;
        const long long prev = G::KB == 3 ? (from1 ? t1 : (have ? -1 : t0)) : t0;
        const u32 prev_n = from1 ? n1 : n0;
        const int pb = from1 ? b1 : b0;
        if (!have && prev < 0) break;
        u32 n_rec = 0;
        u64 old = 0;
        if (have) {
            n_rec = compute(b, tile, have_pre);
            if (t == 0) {
                st_status(&a.tile_status[tile], ((u64)tile_tag4(a) << 32) | n_rec);
                old = atomicAdd(&a.grp_acc[tile / GS4], ACC_ONE4 | (u64)n_rec);
            }
        }
        const u64* bp = prev >= 0 ? tile_base_addr4(a, prev, t) : nullptr;
        u64 bv = 0;
        if (bp != nullptr) bv = ld_status(bp);
        if (have) {
            // did this tile complete its group (the add counts arrivals)? then this warp resolves the group
            const long long g = tile / GS4;
            const long long rest = a.num_tiles - g * GS4;
            const u64 expect = (u64)(rest < GS4 ? rest : GS4);
            const u64 o0 = __shfl_sync(0xffffffffu, old, 0);
            if ((o0 >> 40) == expect - 1) {
                const u64 gsum = (o0 & ACC_MASK4) + n_rec;
                if (t == 0) a.grp_acc[g] = 0;                      // nobody adds to a complete group: clean for the next launch
                const u64 excl = lookback_groups4(a, g, gsum, t);
                if (t == 0 && g == num_groups - 1) {
                    a.total[0] = excl + gsum;                      // pinned host memory: no copy after the kernel
                    if (a.exact_tiles != nullptr) a.total[1] = *reinterpret_cast<volatile unsigned long long*>(a.exact_tiles);
                }
            }
        }
        u64 base = 0;
        if (prev >= 0) base = tile_base_finish4(a, bp, bv, t);
        u32 nx = 0xFFFFFFFFu;
        if (have && t == 0) nx = atomicAdd(a.tile_counter, 1u) - a.ticket_base;
        if (prev >= 0) emit(pb, prev, prev_n, base);
        const long long next_tile = __shfl_sync(0xffffffffu, nx, 0);
        have_pre = 0;
        {
            const long long ng0 = a.org + next_tile * (long long)G::R - G::B_;
            const long long P0 = (ng0 & ~15LL) - a.buf_lo;
            if (next_tile < a.num_tiles && a.vec_ok && P0 >= 0 && P0 + 16LL * G::NWORDS <= a.buf_hi - a.buf_lo) {
                // ONE bulk async copy per tile (cp.async.bulk + mbarrier, issued by lane 0) into the record-slot area of the buffer
                // set the next tile uses: the emit above was the last reader of that area (it ends with __syncwarp), and the bytes
                // are consumed by the next stage before its record slots are written. Nothing is held in registers across the emit
                // (the first version kept three LDG.128 results per lane there: same speed, profiles/r02_zx_sk4_bulk_staging_ab.txt).
                have_pre = 1 | (int)(bulk_phase << 1);
                bulk_phase ^= 1u;
                if (t == 0) {
                    fence_proxy_async_smem();
                    bulk_load(wsm + G::OFF_BUF + ((b + 1) % G::KB) * G::BUF_BYTES + G::BUF_REC, a.bases + P0, (u32)G::RAW_BYTES, mbar);
                }
            }
        }
        // queue: drop what was written, push what was computed
        if (G::KB == 3) {
            if (have || from1) { t1 = t0; n1 = n0; b1 = b0; }
            t0 = have ? tile : -1; n0 = n_rec; b0 = b;
        } else { t0 = have ? tile : -1; n0 = n_rec; b0 = b; }
        tile = next_tile;
        b = (b + 1) % G::KB;
    }
}

__device__ __forceinline__ TileConst4 tile_const4(const SkArgs& a, bool mw)
{
    TileConst4 c;
    c.w = a.w; c.m = a.m; c.k = a.k; c.s8 = (u32)a.s8; c.keymask = a.keymask;
    c.maskhi = mw ? (c.m == 32 ? 0xFFFFFFFFu : ((1u << (2 * c.m - 32)) - 1u)) : 0u;
    c.masklo = mw ? 0xFFFFFFFFu : (c.m == 16 ? 0xFFFFFFFFu : ((1u << (2 * c.m)) - 1u));
    return c;
}

}  // namespace

template <int B, int NB, int Q, int LW, bool CANON, bool MW>
__global__ void __launch_bounds__(SK4_THREADS, Q == 1 ? SK4_MIN_BLOCKS : 4) sk4_kernel(const SkArgs a)
{
    using G = Geo4<B, NB, Q>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int t = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char* wsm = smem_raw + warp * G::WARP_BYTES;
    const TileConst4 c = tile_const4(a, MW);
    tile_loop4<G>(a, wsm, t,
                  [&](int b, long long tile, int have_pre) { return compute_tile4<B, NB, Q, LW, CANON, MW>(a, wsm, c, b, tile, t, have_pre); },
                  [&](int b, long long tile, u32 n_rec, u64 base) { emit_tile4<B, NB, Q, CANON, MW>(a, wsm, c, b, tile, t, n_rec, base); });
}

}  // namespace blg
