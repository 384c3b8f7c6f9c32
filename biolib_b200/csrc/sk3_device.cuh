// sk3_device.cuh -- device code of sk3_kernel.cu / sk3_gen.cu: fused minimizer / super-k-mer extraction for sm_100a, third generation.
//
// Same results as sk2_kernel.cu / sk_kernel.cu (SURVEY.md 8(a) decisions D1-D8; reference intent
// include/minimizer_view.hpp:184-266, include/super_kmer_view.hpp:121-135). What changed is the instruction count:
// the profile of sk2 (profiles/r01_s23_sk2_kernel_instmix.txt) shows ~176 executed instructions per m-mer start with
// the ALU pipe as the busiest unit, 35 of them in the sparse-table window minimum. Here
//
//   * a lane owns a BLOCK of B = (w-1)/Q consecutive m-mer starts (Q = 1 or 2), so the window minimum becomes the
//     van Herk / Gil-Werman scheme entirely in registers: while hashing, the lane keeps the running PREFIX minimum of
//     its block and hands each prefix entry to shared memory; afterwards it walks its block backwards keeping the
//     SUFFIX minimum and combines suffix(i) with the prefix(i) of the lane Q to its right (for Q = 2 the block in
//     between enters as the initial value of the suffix scan). A window [s, s+w) is exactly
//     suffix-of-own-block + (Q-1 whole blocks) + prefix-of-block-(t+Q): three compare/selects per window instead of
//     log2(w)+1, and only the argmin position leaves the last one. Ties: every compare keeps the left operand unless
//     the right one is strictly smaller, so the leftmost minimum wins exactly as the strict '<' of
//     include/minimizer_view.hpp:283,374 does;
//   * each lane cuts its own 64- or 96-bit local copies of the two packed streams (reverse-complement side:
//     complemented little-endian; forward side: big-endian, aligned so that its LAST m-mer sits at bit 0), so every
//     per-m-mer extraction is one or two funnel shifts with immediate amounts, for any tile alignment;
//   * tiles need not start on a 16-base boundary any more: lane 0 is the left margin (the previous window's argmin),
//     lanes 1 .. 31-2Q own the record starts, the Q lanes after them are seen far enough to size those records.
//
// A warp owns a tile end to end and never meets another warp at a barrier: tiles come from a ticket counter, the
// ordered output offset from a two-level chained scan (see tile_base below), and a tile's records are written one tile
// later from the second buffer set, record-parallel with coalesced stores.
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace blg {

namespace {

constexpr int NT = 32;
constexpr int WARPS3 = SK3_THREADS / 32;
constexpr int PADF = 3;     // zero words in front of the big-endian stream

template <int B, int Q>
struct Geo {
    // Q == 0 selects the GENERAL form: any k - m = q B + r with B <= k - m + 1 (q, r and the number of record-owning
    // lanes are launch arguments); Q == 1, 2 are the forms tuned for the benchmark families (r == 0, q == Q)
    static constexpr bool GEN = Q == 0;
    static constexpr int S = NT * B;                   // m-mer starts hashed per tile
    static constexpr int LANES_OUT = 31 - 2 * Q;       // lanes 1 .. 31-2Q own record starts (tuned forms)
    static constexpr int R = LANES_OUT * B;            // new windows per tile (tuned forms)
    static constexpr int NWORDS = S / 16 + 3;          // staged 16-base words: 15 (alignment) + S + 31 (m-1) bases
    static constexpr int NROW = (NWORDS + 8) / 2 + 2;  // u32 words of the packed bad-byte / sequence-start bit rows
    // per-warp shared memory, byte offsets
    static constexpr int OFF_BAD = 0;                              // u32 [NROW]   bit x: base a0 + x is not ACGTU
    static constexpr int OFF_SEQ = OFF_BAD + NROW * 4;             // u32 [NROW]   bit x: a sequence starts at a0 + x
    static constexpr int OFF_FLAG = OFF_SEQ + NROW * 4;            // int [4]
    static constexpr int OFF_P = (OFF_FLAG + 16 + 15) / 16 * 16;   // uint4 [B*32 + Q + 1]  prefix minima, [i][lane]
    static constexpr int OFF_BUF = OFF_P + (B * 32 + (GEN ? 8 : Q + 1)) * 16;
    // double-buffered part (tile i is written out while tile i+1 is computed)
    static constexpr int BUF_CW = 0;                               // u32 [NWORDS + 3]  complemented codes, little-endian
    static constexpr int BUF_EW = BUF_CW + (NWORDS + 3) * 4;       // u32 [PADF + NWORDS + 3]  codes, big-endian words
    static constexpr int BUF_IDXF = (BUF_EW + (PADF + NWORDS + 3) * 4 + 15) / 16 * 16;   // u16 [S]  argmin of every window
    static constexpr int BUF_REC = BUF_IDXF + S * 2;               // u16 [R + 8]  record starts (local window index)
    static constexpr int BUF_EB = (BUF_REC + (R + 8) * 2 + 15) / 16 * 16;                // u32 [S/32 + 4]  record-end bits
    static constexpr int BUF_BYTES = (BUF_EB + (S / 32 + 4) * 4 + 15) / 16 * 16;
    // buffer sets: a tile's records are written out KB-1 tiles after it was computed, which is the slack the chained scan
    // has to deliver the tile's offset without anybody waiting (three sets where shared memory allows 16 warps per SM)
#ifdef SK3_KB
    static constexpr int KB = SK3_KB;
#else
    static constexpr int KB = B <= 12 ? 3 : 2;
#endif
    static constexpr int WARP_BYTES = (OFF_BUF + KB * BUF_BYTES + 127) / 128 * 128;
};

__device__ __forceinline__ u32 pairrev32(u32 x)
{
    x = __brev(x);
    return ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
}

// running (left) keeps unless cand (right) is strictly smaller
__device__ __forceinline__ void take_if_less(u32& rhi, u32& rlo, u32& rix, u32 chi, u32 clo, u32 cix)
{
    const u64 rv = ((u64)rhi << 32) | rlo, cv = ((u64)chi << 32) | clo;
    const bool lt = cv < rv;
    rhi = lt ? chi : rhi; rlo = lt ? clo : rlo; rix = lt ? cix : rix;
}
// cand is to the LEFT of running: it wins unless running is strictly smaller
__device__ __forceinline__ void take_unless_greater(u32& rhi, u32& rlo, u32& rix, u32 chi, u32 clo, u32 cix)
{
    const u64 rv = ((u64)rhi << 32) | rlo, cv = ((u64)chi << 32) | clo;
    const bool keep = rv < cv;
    rhi = keep ? rhi : chi; rlo = keep ? rlo : clo; rix = keep ? rix : cix;
}

// low B bits: OR over d in [0, span) of bit (bitpos + i + d + shift) of the packed bit row, for i = 0 .. B-1
template <int B>
__device__ __forceinline__ u32 or_span_bits(const u32* row, int bitpos, int span, int shift)
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
    return x0 & ((1u << B) - 1u);
}

struct TileConst3 {
    u32 maskhi, masklo;
    int w, m, k;
    u64 s8;
};

template <int B, int Q, int LW, bool CANON, bool MW>
__device__ __forceinline__ u32 compute_tile3(const SkArgs& a, unsigned char* wsm, const TileConst3& c, int b, long long tile, int t,
                                             bool have_pre, uint4 pre0, uint4 pre1)
{
    using G = Geo<B, Q>;
    const int RW = G::GEN ? a.R : G::R;                        // new windows per tile
    const long long g0 = a.org + tile * (long long)RW - B;     // global position of local start 0 (lane 0 = left margin)
    const long long a0 = g0 & ~15LL;                            // staged range starts on a 16-base boundary
    const int o = (int)(g0 - a0);
    unsigned char* buf = wsm + G::OFF_BUF + b * G::BUF_BYTES;
    u32* BAD = reinterpret_cast<u32*>(wsm + G::OFF_BAD);
    u32* SEQ = reinterpret_cast<u32*>(wsm + G::OFF_SEQ);
    int* FLAG = reinterpret_cast<int*>(wsm + G::OFF_FLAG);
    uint4* P = reinterpret_cast<uint4*>(wsm + G::OFF_P);
    u32* CW = reinterpret_cast<u32*>(buf + G::BUF_CW);
    u32* EW = reinterpret_cast<u32*>(buf + G::BUF_EW);
    u32* IDXF32 = reinterpret_cast<u32*>(buf + G::BUF_IDXF);
    unsigned short* REC = reinterpret_cast<unsigned short*>(buf + G::BUF_REC);
    u32* EB = reinterpret_cast<u32*>(buf + G::BUF_EB);

    // ---- stage: ASCII -> packed streams + bad-byte bits -----------------------------------------------------------
    int anybad = 0;
    {
        unsigned short* BAD16 = reinterpret_cast<unsigned short*>(BAD);
        for (int wi = t; wi < 2 * G::NROW; wi += NT) {
            u32 codes = 0, bad = 0xFFFFu;
            if (wi < G::NWORDS) {
                const long long P16 = a0 + 16LL * wi;
                u32 w0, w1, w2, w3;
                if (have_pre) {
                    const uint4 v = wi < NT ? pre0 : pre1;
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
                blg_pack16_bits(w0, w1, w2, w3, &codes, &bad);
                anybad |= (bad != 0);
            }
            if (wi < G::NWORDS + 3) {
                CW[wi] = ~codes;
                EW[PADF + wi] = pairrev32(codes);
            }
            BAD16[wi] = (unsigned short)bad;
        }
        if (t < PADF) EW[t] = 0;
        for (int x = t; x < G::NROW; x += NT) SEQ[x] = 0;
        for (int x = t; x < G::S / 32 + 4; x += NT) EB[x] = 0;
        if (t == 0) FLAG[0] = 0;
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

    // ---- lane-local streams ---------------------------------------------------------------------------------------
    const int x0 = o + B * t;                                  // base index of this lane's first start, from a0
    u32 Rl[LW], Fl[LW];
    {
        const int wi0 = x0 >> 4, sh = (x0 & 15) * 2;
        u32 cw[LW + 1];
#pragma unroll
        for (int j = 0; j <= LW; ++j) cw[j] = CW[wi0 + j];
#pragma unroll
        for (int j = 0; j < LW; ++j) Rl[j] = __funnelshift_r(cw[j], cw[j + 1], sh);   // Rl[0] = least significant
        // big-endian side: the 32*LW bits that END with the last base of this lane's last m-mer
        const int bp = 2 * (x0 + B - 1 + c.m) - 32 * LW + 32 * PADF;
        const int wj = bp >> 5, fs = bp & 31;
        u32 ew[LW + 1];
#pragma unroll
        for (int j = 0; j <= LW; ++j) ew[j] = EW[wj + j];
#pragma unroll
        for (int j = 0; j < LW; ++j) Fl[j] = __funnelshift_l(ew[j + 1], ew[j], fs);   // Fl[0] = most significant
    }

    // ---- one hash per m-mer start, running prefix minimum of the block ----------------------------------------------
    u32 hh[B], hl[B];
    {
        u32 phi = 0, plo = 0, pix = 0;
#pragma unroll
        for (int i = 0; i < B; ++i) {
            const int cI = 2 * (B - 1 - i);
            u32 flo, fhi = 0;
            if (LW == 2) {
                flo = __funnelshift_r(Fl[1], Fl[0], cI);
                if (MW) fhi = (Fl[0] >> cI) & c.maskhi; else flo &= c.masklo;
            } else {
                flo = __funnelshift_r(Fl[2], Fl[1], cI);
                if (MW) fhi = __funnelshift_r(Fl[1], Fl[0], cI) & c.maskhi; else flo &= c.masklo;
            }
            if (G::GEN) flo &= c.masklo;            // general form: any m, both words masked
            u64 val = ((u64)fhi << 32) | flo;
            if (CANON) {
                u32 rlo = __funnelshift_r(Rl[0], Rl[1], 2 * i), rhi = 0;
                if (MW) rhi = (LW == 2 ? (Rl[1] >> (2 * i)) : __funnelshift_r(Rl[1], Rl[LW - 1], 2 * i)) & c.maskhi;
                else rlo &= c.masklo;
                if (G::GEN) rlo &= c.masklo;
                const u64 rc = ((u64)rhi << 32) | rlo;
                val = rc < val ? rc : val;
            }
            const u64 h = murmur3_key8(val, c.s8);
            hh[i] = (u32)(h >> 32); hl[i] = (u32)h;
            if (i == 0) { phi = hh[0]; plo = hl[0]; pix = (u32)(B * t); }
            else take_if_less(phi, plo, pix, hh[i], hl[i], (u32)(B * t + i));
            P[i * 32 + t] = make_uint4(phi, plo, pix, 0u);
        }
    }
    __syncwarp();

    // ---- suffix minimum + combine: argmin position of every window ----------------------------------------------------
    u32 vix[B];
    {
        u32 shi = 0, slo = 0, six = 0;
        if (G::GEN) {
            // k - m = q B + r: window (t, i) = own suffix from i + whole blocks t+1 .. t+q-1 (seed of the scan) + block t+q
            // when i + r >= B (and q >= 1) + prefix of block t + q + [i + r >= B] through index (i + r) mod B
            const int q = a.gq, r = a.gr;
            if (q >= 2) {
                const uint4 e = P[(B - 1) * 32 + t + 1]; shi = e.x; slo = e.y; six = e.z;
                for (int x = 2; x < q; ++x) { const uint4 f = P[(B - 1) * 32 + t + x]; take_if_less(shi, slo, six, f.x, f.y, f.z); }
            }
            const uint4 bq = P[(B - 1) * 32 + t + (q >= 1 ? q : 0)];
            // prefix entries: row (i + r) mod B of block t + q (+1 after the wrap): two base pointers, the row offset
            // i * 32 is then an immediate
            const uint4* pa = P + r * 32 + t + q;
            const uint4* pb = pa - B * 32 + 1;
#pragma unroll
            for (int i = B - 1; i >= 0; --i) {
                if (i == B - 1 && q < 2) { shi = hh[i]; slo = hl[i]; six = (u32)(B * t + i); }
                else take_unless_greater(shi, slo, six, hh[i], hl[i], (u32)(B * t + i));
                const bool wrap = i + r >= B;
                u32 chi = shi, clo = slo, cix = six;
                if (wrap && q >= 1) take_if_less(chi, clo, cix, bq.x, bq.y, bq.z);
                const uint4 e = (wrap ? pb : pa)[i * 32];
                const u64 sv = ((u64)chi << 32) | clo, ev = ((u64)e.x << 32) | e.y;
                vix[i] = ev < sv ? e.z : cix;
            }
        } else {
            if (Q == 2) { const uint4 e = P[(B - 1) * 32 + t + 1]; shi = e.x; slo = e.y; six = e.z; }
#pragma unroll
            for (int i = B - 1; i >= 0; --i) {
                if (Q == 1 && i == B - 1) { shi = hh[i]; slo = hl[i]; six = (u32)(B * t + i); }
                else take_unless_greater(shi, slo, six, hh[i], hl[i], (u32)(B * t + i));
                const uint4 e = P[i * 32 + t + Q];
                const u64 sv = ((u64)shi << 32) | slo, ev = ((u64)e.x << 32) | e.y;
                vix[i] = ev < sv ? e.z : six;
            }
        }
    }
#pragma unroll
    for (int i = 0; i < B; i += 2) IDXF32[(B / 2) * t + (i >> 1)] = vix[i] | (vix[i + 1] << 16);

    // ---- validity, boundaries -----------------------------------------------------------------------------------------
    constexpr u32 MB = (1u << B) - 1u;
    u32 V = MB;
    if (dirty | seqdirty) {
        u32 inv = 0;
        if (dirty) inv |= or_span_bits<B>(BAD, x0, c.k, 0);
        if (seqdirty && c.k > 1) inv |= or_span_bits<B>(SEQ, x0, c.k - 1, 1);
        V = ~inv & MB;
    }
    u32 dif = 0;
#pragma unroll
    for (int i = 0; i + 1 < B; ++i) dif |= (u32)(vix[i + 1] != vix[i]) << i;
    {
        const u32 nv0 = __shfl_down_sync(0xffffffffu, vix[0], 1);
        dif |= (u32)(nv0 != vix[B - 1]) << (B - 1);
    }
    const u32 Vn = (V >> 1) | ((__shfl_down_sync(0xffffffffu, V, 1) & 1u) << (B - 1));
    const u32 E = V & (~Vn | dif) & MB;        // the record ends with this window
    const u32 X = (E | ~V) & MB;               // the next window cannot continue a record
    {
        const int bp = B * t, wq = bp >> 5, sh = bp & 31;
        if (E) {
            atomicOr(&EB[wq], E << sh);
            if (sh + B > 32) atomicOr(&EB[wq + 1], E >> (32 - sh));
        }
    }
    const u32 xprev = (__shfl_up_sync(0xffffffffu, X, 1) >> (B - 1)) & 1u;

    // ---- record starts ---------------------------------------------------------------------------------------------------
    u32 start = 0;
    if (t >= 1 && t <= (G::GEN ? a.gl : G::LANES_OUT)) {
        start = V & ((X << 1) | xprev) & MB;
        // windows outside [win_lo, win_hi) belong to another call / rank
        const long long gl = g0 + (long long)B * t;
        long long nlo = a.win_lo - gl, nhi = a.win_hi - gl;
        nlo = nlo < 0 ? 0 : (nlo > B ? B : nlo);
        nhi = nhi < 0 ? 0 : (nhi > B ? B : nhi);
        start &= (MB << nlo) & ((1u << nhi) - 1u);
    }
    const u32 cnt = __popc(start);
    u32 incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 v = __shfl_up_sync(0xffffffffu, incl, d);
        if (t >= d) incl += v;
    }
    u32 off = incl - cnt, sm = start;
    while (sm) {
        const int i = __ffs((int)sm) - 1;
        sm &= sm - 1;
        REC[off++] = (unsigned short)(B * t + i);
    }
    __syncwarp();
    return __shfl_sync(0xffffffffu, incl, 31);
}

// Record-parallel output of a finished tile from buffer set b: `n_rec` records, the first one at global index `base`.
template <int B, int Q, int LW, bool CANON, bool MW>
__device__ __forceinline__ void emit_tile3(const SkArgs& a, const unsigned char* wsm, const TileConst3& c, int b, long long tile, int t,
                                           u32 n_rec, u64 base)
{
    using G = Geo<B, Q>;
    const long long g0 = a.org + tile * (long long)(G::GEN ? a.R : G::R) - B;
    const int o = (int)(g0 - (g0 & ~15LL));
    const unsigned char* buf = wsm + G::OFF_BUF + b * G::BUF_BYTES;
    const u32* CW = reinterpret_cast<const u32*>(buf + G::BUF_CW);
    const u32* EW = reinterpret_cast<const u32*>(buf + G::BUF_EW);
    const unsigned short* IDXF = reinterpret_cast<const unsigned short*>(buf + G::BUF_IDXF);
    const unsigned short* REC = reinterpret_cast<const unsigned short*>(buf + G::BUF_REC);
    const u32* EB = reinterpret_cast<const u32*>(buf + G::BUF_EB);
    for (u32 r = t; r < n_rec; r += NT) {
        const int s0 = REC[r];
        const int j = IDXF[s0];
        // size: distance to the first record-end bit at or after s0, at most w
        int size;
        {
            const int wq = s0 >> 5, sh = s0 & 31;
            const u32 e0 = EB[wq], e1 = EB[wq + 1];
            u32 bits = __funnelshift_r(e0, e1, sh);
            if (bits) size = __ffs((int)bits);
            else {
                bits = __funnelshift_r(e1, EB[wq + 2], sh);
                size = bits ? 32 + __ffs((int)bits) : c.w;
            }
            size = size < c.w ? size : c.w;
        }
        const int xj = o + j;
        u64 val;
        {
            const int e = 2 * xj + 2 * c.m + 32 * PADF - 64;     // the 64 bits that end with the m-mer's last base
            const int wj = e >> 5, fs = e & 31;
            const u32 e0 = EW[wj], e1 = EW[wj + 1], e2 = EW[wj + 2];
            u32 fhi = __funnelshift_l(e1, e0, fs), flo = __funnelshift_l(e2, e1, fs);
            if (MW) fhi &= c.maskhi; else { fhi = 0; flo &= c.masklo; }
            if (G::GEN) flo &= c.masklo;
            val = ((u64)fhi << 32) | flo;
        }
        if (CANON) {
            const int wi = xj >> 4, sh = (xj & 15) * 2;
            const u32 c0 = CW[wi], c1 = CW[wi + 1], c2 = CW[wi + 2];
            u32 rlo = __funnelshift_r(c0, c1, sh), rhi = __funnelshift_r(c1, c2, sh);
            if (MW) rhi &= c.maskhi; else { rhi = 0; rlo &= c.masklo; }
            if (G::GEN) rlo &= c.masklo;
            const u64 rc = ((u64)rhi << 32) | rlo;
            val = rc < val ? rc : val;
        }
        const u64 g = base + r;
        if (g < a.capacity) {
            a.value[g] = val;
            store_mm_start(a, g, g0 + j);
            if (a.size != nullptr) { a.size[g] = (u8)size; a.mm_pos[g] = (u8)(j - s0); }
            if (a.hash != nullptr) a.hash[g] = murmur3_key8(val, c.s8);
        }
    }
    __syncwarp();   // this buffer set is restaged by the next compute_tile
}

// ---- ordered output offset without any CTA barrier: a two-level chained scan ------------------------------------------
// Tiles are handed out by ticket, one per warp. Level 1: a tile publishes its record count (one store) and adds it to
// the accumulator of its GROUP of 32 consecutive tiles (one 64-bit atomicAdd that also counts arrivals). Level 2: the
// warp whose add completes a group publishes the group total and resolves the group's prefix with a decoupled look-back
// over GROUPS (32x fewer words than tiles, so one round of 32 loads reaches an inclusive word). A tile's base is then
// exclusive(group) + the counts of the earlier tiles of its group: one load per lane, issued one tile after everything
// it reads was written. Nothing here waits on a warp holding a HIGHER ticket, so the spins cannot deadlock.
constexpr int GS = 32;
constexpr u64 TS_SET = 1ULL << 62;
constexpr u64 ACC_ONE = 1ULL << 40, ACC_MASK = ACC_ONE - 1;

__device__ __forceinline__ void spin_guard(u32& spins)
{
    __nanosleep(64);
#ifdef BLG_DEBUG_SPIN
    if (++spins > (1u << 28)) __trap();   // debug builds only (as in sk4): a protocol bug must not hang the GPU; a release build
#else                                     // never traps, because under a debugger, sanitizer or time-slicing a legitimate wait can be long
    (void)spins;
#endif
}

// Exclusive record offset of `tile` (published one tile ago), in two steps so that the loads are in flight while the
// warp does its group duty: lanes below the tile's index in its group read the earlier tiles' counts, lane 31 the
// inclusive prefix of the previous group.
__device__ __forceinline__ const u64* tile_base_addr(const SkArgs& a, long long tile, int t)
{
    const long long g = tile / GS;
    const int idx = (int)(tile - g * GS);
    if (t < idx) return a.tile_status + g * GS + t;
    if (t == 31 && g > 0) return a.grp_status + (g - 1);
    return nullptr;
}
__device__ __forceinline__ u64 tile_base_finish(const u64* p, u64 v, int t)
{
    if (p != nullptr) {
        u32 spins = 0;
        if (t == 31) { while ((v >> 62) != 2) { spin_guard(spins); v = ld_status(p); } }
        else { while (!(v & TS_SET)) { spin_guard(spins); v = ld_status(p); } }
        v &= ST_VAL;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

}  // namespace

template <int B, int Q, int LW, bool CANON, bool MW>
__global__ void __launch_bounds__(SK3_THREADS, SK3_MIN_BLOCKS) sk3_kernel(const SkArgs a)
{
    using G = Geo<B, Q>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int t = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char* wsm = smem_raw + warp * G::WARP_BYTES;

    TileConst3 c;
    c.w = a.w; c.m = a.m; c.k = a.k; c.s8 = a.s8;
    if (G::GEN) {
        c.maskhi = c.m > 16 ? (c.m == 32 ? 0xFFFFFFFFu : ((1u << (2 * c.m - 32)) - 1u)) : 0u;
        c.masklo = c.m >= 16 ? 0xFFFFFFFFu : ((1u << (2 * c.m)) - 1u);
    } else {
        c.maskhi = MW ? (c.m == 32 ? 0xFFFFFFFFu : ((1u << (2 * c.m - 32)) - 1u)) : 0u;
        c.masklo = MW ? 0xFFFFFFFFu : (c.m == 16 ? 0xFFFFFFFFu : ((1u << (2 * c.m)) - 1u));
    }

    const long long num_groups = (a.num_tiles + GS - 1) / GS;
    u32 tk = 0;
    if (t == 0) tk = atomicAdd(a.tile_counter, 1u);
    long long tile = __shfl_sync(0xffffffffu, tk, 0);
    // computed, not yet written out: (t0, n0, b0) the newer, (t1, n1, b1) the older (KB == 3 only): tile, records, buffer set
    long long t0 = -1, t1 = -1;
    u32 n0 = 0, n1 = 0;
    int b0 = 0, b1 = 0;
    bool have_pre = false;
    uint4 pre0 = make_uint4(0, 0, 0, 0), pre1 = pre0;
    int b = 0;                          // buffer set of the tile computed in this iteration
    // Every warp runs on its own: compute tile i, publish its count, write out tile i-(KB-1) from its buffer set.
    // A ticket is taken only when nothing can block any more before that tile is published (a warp that waits for a
    // prefix while it holds an uncomputed tile stalls that tile's group, the groups behind it, and so itself: convoys).
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
            n_rec = compute_tile3<B, Q, LW, CANON, MW>(a, wsm, c, b, tile, t, have_pre, pre0, pre1);
            if (t == 0) {
                st_status(&a.tile_status[tile], TS_SET | n_rec);
                old = atomicAdd(&a.grp_acc[tile / GS], ACC_ONE | (u64)n_rec);
            }
        }
        const u64* bp = prev >= 0 ? tile_base_addr(a, prev, t) : nullptr;
        u64 bv = 0;
        if (bp != nullptr) bv = ld_status(bp);
        if (have) {
            // did this tile complete its group (the add counts arrivals)? then this warp resolves the group
            const long long g = tile / GS;
            const long long rest = a.num_tiles - g * GS;
            const u64 expect = (u64)(rest < GS ? rest : GS);
            const u64 o0 = __shfl_sync(0xffffffffu, old, 0);
            if ((o0 >> 40) == expect - 1) {
                const u64 gsum = (o0 & ACC_MASK) + n_rec;
                const u64 excl = lookback_exclusive(a.grp_status, g, gsum, t);
                if (t == 0 && g == num_groups - 1) *a.total = excl + gsum;
            }
        }
        u64 base = 0;
        if (prev >= 0) base = tile_base_finish(bp, bv, t);
        u32 nx = 0xFFFFFFFFu;
        if (have && t == 0) nx = atomicAdd(a.tile_counter, 1u);
        if (prev >= 0) emit_tile3<B, Q, LW, CANON, MW>(a, wsm, c, pb, prev, t, prev_n, base);
        const long long next_tile = __shfl_sync(0xffffffffu, nx, 0);
        have_pre = false;
        {
            const long long ng0 = a.org + next_tile * (long long)(G::GEN ? a.R : G::R) - B;
            const long long P0 = (ng0 & ~15LL) - a.buf_lo;
            if (next_tile < a.num_tiles && a.vec_ok && P0 >= 0 && P0 + 16LL * G::NWORDS <= a.buf_hi - a.buf_lo) {
                have_pre = true;
                if (t < G::NWORDS) pre0 = __ldg(reinterpret_cast<const uint4*>(a.bases + P0) + t);
                if (G::NWORDS > NT && t < G::NWORDS - NT) pre1 = __ldg(reinterpret_cast<const uint4*>(a.bases + P0) + NT + t);
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

}  // namespace blg
