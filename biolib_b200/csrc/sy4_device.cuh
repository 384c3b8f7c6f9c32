// sy4_device.cuh -- closed syncmers (sampler::syncmer_sampler over hash::minimizer_position_extractor, reference
// include/syncmer_sampler.hpp:130-137, include/kmer_view.hpp:266-283) in the frame of the sk4 kernel, for k - m = 12 or 10.
//
// A k-mer is kept iff the offset of its minimum-hash m-mer is sync_s or sync_e. The extractor walks the bit fields of the k-mer
// VALUE: for a forward value these are the forward m-mers of starts p .. p + w - 1 (offset = start - p, leftmost minimum wins:
// it scans from the right with '>='); for a reverse-complement value (canonical input, rc < fwd) they are the reverse-complement
// m-mers of the same starts in reverse order (offset = p + w - 1 - start, so the RIGHTMOST minimum in forward coordinates wins).
// The first fused filter (kmer2_kernel<.., SYNC>) hashed every m-mer once per strand and let each k-mer walk its w hashes:
// 460 instructions per k-mer. Here the window minimum is sk4's: one 32-bit key per m-mer start and strand (hash bits above an
// index field), prefix / suffix minima over blocks of B = w - 1 starts in registers, a hash-part tie caught by the key
// differences and redone exactly. Two rows:
//   F  forward m-mers, index field i            (smaller index wins a tie; keys of the next block enter with +16)
//   R  reverse-complement m-mers, field 15 - i  (larger index wins a tie; keys of the OWN block enter with +16)
// The argmin bytes of both rows go to shared memory; the keep bits of a lane's windows come out of them four windows at a
// time; the strand of a k-mer is decided by the top 16 bases of its two values (full compare when those are equal).
#pragma once
#include "sk4_device.cuh"

namespace blg {

namespace {

struct SyncConst4 {
    u32 kmaskhi, kmasklo;      // masks of a k-mer value's two words
    u32 tF0, tF1, tR0, tR1;    // wanted minimum offsets (forward coordinates), replicated into four bytes: F row s, e; R row w-1-s, w-1-e
};

// forward / reverse-complement value of the `len`-mer at local base index xj (len <= 32), masks given
__device__ __forceinline__ void strands_from_smem(const u32* CW, const u32* EW, int xj, int len, u32 maskhi, u32 masklo, u64* f, u64* r)
{
    {
        const int e = 2 * xj + 2 * len + 32 * PADF4 - 64;
        const int wj = e >> 5, fs = e & 31;
        const u32 e0 = EW[wj], e1 = EW[wj + 1], e2 = EW[wj + 2];
        const u32 fhi = __funnelshift_l(e1, e0, fs) & maskhi, flo = __funnelshift_l(e2, e1, fs) & masklo;
        *f = ((u64)fhi << 32) | flo;
    }
    {
        const int wi = xj >> 4, sh = (xj & 15) * 2;
        const u32 c0 = CW[wi], c1 = CW[wi + 1], c2 = CW[wi + 2];
        const u32 rlo = __funnelshift_r(c0, c1, sh) & masklo, rhi = __funnelshift_r(c1, c2, sh) & maskhi;
        *r = ((u64)rhi << 32) | rlo;
    }
}

// windows of one block of one row: own keys kk[] (suffix side) and the prefix keys np[] of the next block -> argmin bytes
template <int B, bool REV>
__device__ __forceinline__ void finish_sync_block4(const u32 (&kk)[B], const u32 (&np)[(B + 3) / 4 * 4], u32& tacc, unsigned char* dst)
{
    u32 mvb[B];
    u32 sm = 0;
#pragma unroll
    for (int i = B - 1; i >= 0; --i) {
        if (i == B - 1) sm = kk[i];
        else {
            tacc = min(tacc, REV ? kk[i] - sm : sm - kk[i]);     // equal hash parts: the index fields differ by a small number >= 0
            sm = min(kk[i], sm);
        }
        tacc = min(tacc, REV ? np[i] - sm : sm - np[i]);
        mvb[i] = REV ? min(np[i], sm + 16u) : min(np[i] + 16u, sm);
    }
    u32 wv[4] = {0, 0, 0, 0};
#pragma unroll
    for (int q = 0; q < (B + 3) / 4; ++q) {
        const u32 m0 = mvb[4 * q], m1 = 4 * q + 1 < B ? mvb[4 * q + 1] : 0u, m2 = 4 * q + 2 < B ? mvb[4 * q + 2] : 0u,
                  m3 = 4 * q + 3 < B ? mvb[4 * q + 3] : 0u;
        wv[q] = __byte_perm(__byte_perm(m0, m1, 0x0040), __byte_perm(m2, m3, 0x0040), 0x5410);
    }
    *reinterpret_cast<uint4*>(dst) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
}

// One row (strand) of a tile: every lane hashes the m-mers of its NB blocks and leaves the argmin byte of each of its windows in
// IDX[lane][block][window]. F row byte: bits 0-3 index in its block, bit 4 = 1 if in the NEXT block. R row byte: bits 0-3 =
// 15 - index, bit 4 = 1 if in the OWN block.
template <int B, int NB, int NP, int LW, bool REV, bool MW>
__device__ __forceinline__ void sync_row4(const u32* CW, const u32* EW, uint4* P, unsigned char* IDX, const TileConst4& c, int x0, int t, u32& tacc)
{
    constexpr int BQ = (B + 3) / 4;
    u32 ks_prev[B];
#pragma unroll 1
    for (int nb = 0; nb < NB; ++nb) {
        const int xb = x0 + nb * B;
        u32 W[LW];                                         // block-local stream of this strand
        if (REV) {
            const int wi0 = xb >> 4, sh = (xb & 15) * 2;
            u32 cw[LW + 1];
#pragma unroll
            for (int j = 0; j <= LW; ++j) cw[j] = CW[wi0 + j];
#pragma unroll
            for (int j = 0; j < LW; ++j) W[j] = __funnelshift_r(cw[j], cw[j + 1], sh);        // W[0] = least significant
        } else {
            const int bp = 2 * (xb + B - 1 + c.m) - 32 * LW + 32 * PADF4;
            const int wj = bp >> 5, fs = bp & 31;
            u32 ew[LW + 1];
#pragma unroll
            for (int j = 0; j <= LW; ++j) ew[j] = EW[wj + j];
#pragma unroll
            for (int j = 0; j < LW; ++j) W[j] = __funnelshift_l(ew[j + 1], ew[j], fs);        // W[0] = most significant
        }
        u32 ks[B];
        u32 pk[BQ * 4];
        u32 pm = 0;
#pragma unroll
        for (int i = 0; i < B; ++i) {
            u32 lo, hi = 0;
            if (REV) {
                const int bo = 2 * i, wq = bo >> 5, sh = bo & 31;
                lo = __funnelshift_r(word_or_zero(W, wq), word_or_zero(W, wq + 1), sh);
                if (MW) hi = __funnelshift_r(word_or_zero(W, wq + 1), word_or_zero(W, wq + 2), sh) & c.maskhi;
                else lo &= c.masklo;
            } else {
                const int cI = 2 * (B - 1 - i), wq = cI >> 5, sh = cI & 31;
                lo = __funnelshift_r(word_or_zero(W, LW - 1 - wq), word_or_zero(W, LW - 2 - wq), sh);
                if (MW) hi = __funnelshift_r(word_or_zero(W, LW - 2 - wq), word_or_zero(W, LW - 3 - wq), sh) & c.maskhi;
                else lo &= c.masklo;
            }
            const u32 hh = (u32)(murmur3_key8_s32(((u64)hi << 32) | lo, c.s8) >> 32);
            ks[i] = make_key(hh, c.keymask, REV ? (u32)(15 - i) : (u32)i);
            if (i == 0) pm = ks[i];
            else {
                tacc = min(tacc, REV ? pm - ks[i] : ks[i] - pm);
                pm = min(pm, ks[i]);
            }
            pk[i] = pm;
        }
#pragma unroll
        for (int i = B; i < BQ * 4; ++i) pk[i] = 0;
        if (nb == 0) {
#pragma unroll
            for (int q = 0; q < BQ; ++q) P[q * 32 + t] = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
        }
        if (nb >= 1) finish_sync_block4<B, REV>(ks_prev, pk, tacc, IDX + t * 32 + (nb - 1) * 16);
#pragma unroll
        for (int i = 0; i < B; ++i) ks_prev[i] = ks[i];
    }
    __syncwarp();
    {
        u32 np[BQ * 4];                                    // the last block: its next block is the next lane's block 0
#pragma unroll
        for (int q = 0; q < BQ; ++q) {
            const uint4 e = P[q * 32 + t + 1];
            np[4 * q] = e.x; np[4 * q + 1] = e.y; np[4 * q + 2] = e.z; np[4 * q + 3] = e.w;
        }
        finish_sync_block4<B, REV>(ks_prev, np, tacc, IDX + t * 32 + (NB - 1) * 16);
    }
}

// Exact redo of one row for the windows a tie could have changed (see redo_windows4): full 64-bit hashes through shuffles, a strict
// '<' (leftmost) for the F row, '<=' (rightmost) for the R row; rewrites the argmin bytes in the row's format.
template <int B, int NB, bool REV, bool MW>
__device__ __noinline__ void redo_sync4(const u32* CW, const u32* EW, unsigned char* IDX, TileConst4 c, int o, int t, u32 flagged)
{
    constexpr int L = NB * B, W = B + 1, S = 32 * L;
    constexpr int NU = 32 - W + 1, UNITS = (S + NU - 1) / NU;
    static_assert(W <= 17 && UNITS <= 64, "one hashed start per lane and unit; the unit mask is 64 bits");
    u64 todo = 0;
    for (u32 fm = flagged; fm; fm &= fm - 1) {
        const int f = __ffs((int)fm) - 1;
        int lo = L * f - B, hi = L * f + L;
        lo = lo < 0 ? 0 : lo;
        const int u0 = lo / NU, u1 = (hi - 1) / NU;
        for (int u = u0; u <= u1 && u < UNITS; ++u) todo |= 1ULL << u;
    }
    for (; todo; todo &= todo - 1) {
        const int s = (__ffsll((long long)todo) - 1) * NU;
        u64 f, r;
        strands_from_smem(CW, EW, o + s + t, c.m, MW ? c.maskhi : 0u, c.masklo, &f, &r);
        const u64 h = murmur3_key8_s32(REV ? r : f, c.s8);
        u64 best = h;
        int bd = 0;
#pragma unroll 4
        for (int d = 1; d < W; ++d) {
            const u64 hd = __shfl_sync(0xffffffffu, h, (t + d) & 31);
            if (REV ? hd <= best : hd < best) { best = hd; bd = d; }
        }
        const int ws = s + t;
        if (t < NU && ws < S) {
            const int gb = ws / B, i = ws - gb * B;
            const int pb = i + bd, pv = pb >= B ? 1 : 0, idx = pb - pv * B;
            IDX[gb * 16 + i] = (unsigned char)(REV ? (((1 - pv) << 4) | (15 - idx)) : ((pv << 4) | idx));
        }
    }
    __syncwarp();
}

// bit j: the minimum of the lane's window j sits at one of the two wanted offsets (tA / tB: the offset in forward coordinates in
// all four bytes). Works on the packed argmin bytes: byte -> position relative to the block's first start -> minus the window's
// own index -> compared with both targets, four windows per step.
template <int B, int NB, bool REV>
__device__ __forceinline__ u32 keep_bits4(const unsigned char* IDX, int t, u32 tA, u32 tB)
{
    const uint4* IW = reinterpret_cast<const uint4*>(IDX + t * 32);
    u32 keep = 0;
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
        const uint4 e = IW[nb];
        const u32 wv[4] = {e.x, e.y, e.z, e.w};
#pragma unroll
        for (int q = 0; q < (B + 3) / 4; ++q) {
            const u32 lo = wv[q] & 0x0F0F0F0Fu, hi = (wv[q] >> 4) & 0x01010101u;
            // F: pos = idx + B * next;  R: pos = (15 - field) + B * (1 - own)
            const u32 pos = REV ? (0x0F0F0F0Fu + (u32)B * 0x01010101u) - lo - hi * (u32)B : lo + hi * (u32)B;
            const u32 d = pos - (0x03020100u + 0x04040404u * (u32)q);              // minus the window's index in its block: 0 .. B per byte
            const u32 za = ((d ^ tA) + 0x7F7F7F7Fu) & 0x80808080u, zb = ((d ^ tB) + 0x7F7F7F7Fu) & 0x80808080u;   // 0x80 where the byte differs
            const u32 eq = ~(za & zb) & 0x80808080u;                                 // equal to tA or to tB
            const u32 NIB = B - 4 * q >= 4 ? 0xFu : ((1u << ((B - 4 * q) & 3)) - 1u);     // the last word of a block may hold padding
            keep |= (((eq * 0x00204081u) >> 28) & NIB) << (nb * B + 4 * q);
        }
    }
    constexpr int L = NB * B;
    return keep & (L >= 32 ? 0xFFFFFFFFu : ((1u << (L & 31)) - 1u));
}

// for the windows in `need` (bit j: the lane's window j): bit j set iff the reverse complement of that k-mer is smaller than the
// forward k-mer (the canonical value is the reverse complement). The top 16 bases of both decide unless they are equal.
__device__ __forceinline__ u32 strand_bits4(const u32* CW, const u32* EW, const TileConst4& c, const SyncConst4& sc, int x0, u32 need)
{
    u32 sr = 0;
    while (need) {
        const int j = __ffs((int)need) - 1;
        need &= need - 1;
        const int xj = x0 + j;
        bool use_r, decided = false;
        if (c.k >= 16) {
            const int e = 2 * xj + 32 * PADF4, wj = e >> 5, fs = e & 31;
            const u32 ftop = __funnelshift_l(EW[wj + 1], EW[wj], fs);                 // bases xj .. xj+15, the first one on top
            const int y = xj + c.k - 16, wi = y >> 4, sh = (y & 15) * 2;
            const u32 rtop = __funnelshift_r(CW[wi], CW[wi + 1], sh);                 // complement of bases xj+k-16 .. xj+k-1, the last one on top
            use_r = rtop < ftop;
            decided = rtop != ftop;
        }
        if (!decided) {
            u64 f, r;
            strands_from_smem(CW, EW, xj, c.k, sc.kmaskhi, sc.kmasklo, &f, &r);
            use_r = r < f;
        }
        sr |= (u32)use_r << j;
    }
    return sr;
}

// Stage + both rows + keep bits + slots of the kept k-mers of one tile into buffer set b. Returns the number of kept k-mers.
template <int B, int NB, int LW, bool CANON, bool MW>
__device__ __forceinline__ u32 compute_sync4(const SkArgs& a, unsigned char* wsm, const TileConst4& c, const SyncConst4& sc, int b, long long tile,
                                             int t, int have_pre)
{
    constexpr int NP = CANON ? 2 : 1;
    using G = Geo4<B, NB, 1, NP>;
    constexpr int L = G::L;
    constexpr u32 ML = L >= 32 ? 0xFFFFFFFFu : ((1u << (L & 31)) - 1u);
    const long long g0 = a.org + tile * (long long)G::R - B;
    int o;
    const u32 V = stage_tile4<B, NB, 1, NP>(a, wsm, c, b, tile, t, have_pre, &o);
    unsigned char* buf = wsm + G::OFF_BUF + b * G::BUF_BYTES;
    uint4* PF = reinterpret_cast<uint4*>(wsm + G::OFF_P);
    uint4* PR = PF + (G::BQ * 32 + 1);
    const u32* CW = reinterpret_cast<const u32*>(buf + G::BUF_CW);
    const u32* EW = reinterpret_cast<const u32*>(buf + G::BUF_EW);
    unsigned char* IDXF = buf + G::BUF_IDX;
    unsigned char* IDXR = wsm + G::OFF_IDX2;
    unsigned short* REC = reinterpret_cast<unsigned short*>(buf + G::BUF_REC);
    const int x0 = o + L * t;

    u32 tacc = 0xFFFFFFFFu;
    sync_row4<B, NB, NP, LW, false, MW>(CW, EW, PF, IDXF, c, x0, t, tacc);
    if (CANON) sync_row4<B, NB, NP, LW, true, MW>(CW, EW, PR, IDXR, c, x0, t, tacc);
    const u32 flagged = __ballot_sync(0xffffffffu, tacc < 32u && V != 0u);
    if (flagged) {
        __syncwarp();
        redo_sync4<B, NB, false, MW>(CW, EW, IDXF, c, o, t, flagged);
        if (CANON) redo_sync4<B, NB, true, MW>(CW, EW, IDXR, c, o, t, flagged);
        if (t == 0 && a.exact_tiles != nullptr) atomicAdd(a.exact_tiles, 1ULL);
    }
    u32 keep = keep_bits4<B, NB, false>(IDXF, t, sc.tF0, sc.tF1);
    if (CANON) {
        const u32 kr = keep_bits4<B, NB, true>(IDXR, t, sc.tR0, sc.tR1);
        const u32 sr = strand_bits4(CW, EW, c, sc, x0, (keep ^ kr) & V);       // the strand only matters where the rows disagree
        keep = (keep & ~sr) | (kr & sr);
    }
    u32 start = keep & V & ML;
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

// the kept k-mers of a finished tile: (value, position) in tile order
template <int B, int NB, bool CANON>
__device__ __forceinline__ void emit_sync4(const SkArgs& a, const unsigned char* wsm, const TileConst4& c, const SyncConst4& sc, int b, long long tile,
                                           int t, u32 n_rec, u64 base)
{
    using G = Geo4<B, NB, 1, CANON ? 2 : 1>;
    constexpr int L = G::L;
    const long long g0 = a.org + tile * (long long)G::R - B;
    const int o = (int)(g0 - (g0 & ~15LL));
    const unsigned char* buf = wsm + G::OFF_BUF + b * G::BUF_BYTES;
    const u32* CW = reinterpret_cast<const u32*>(buf + G::BUF_CW);
    const u32* EW = reinterpret_cast<const u32*>(buf + G::BUF_EW);
    const unsigned short* REC = reinterpret_cast<const unsigned short*>(buf + G::BUF_REC);
    for (u32 r = t; r < n_rec; r += NT4) {
        const u32 rec = REC[r];
        const int j = L * (int)(rec >> 5) + (int)((rec >> 4) & 1u) * B + (int)(rec & 15u);   // local k-mer start
        u64 f, rc;
        strands_from_smem(CW, EW, o + j, c.k, sc.kmaskhi, sc.kmasklo, &f, &rc);
        const u64 g = base + r;
        if (g < a.capacity) {
            a.value[g] = CANON && rc < f ? rc : f;
            if (a.mm_start != nullptr) a.mm_start[g] = (unsigned long long)(g0 + j);
        }
    }
    __syncwarp();
}

}  // namespace

// forward-only input: one row, 96 registers, 5 CTAs per SM as sk4; canonical input: two rows (and their shared memory), 4 CTAs
template <int B, int NB, int LW, bool CANON, bool MW>
__global__ void __launch_bounds__(SK4_THREADS, CANON ? 4 : SK4_MIN_BLOCKS) sy4_kernel(const SkArgs a)
{
    using G = Geo4<B, NB, 1, CANON ? 2 : 1>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int t = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char* wsm = smem_raw + warp * G::WARP_BYTES;
    const TileConst4 c = tile_const4(a, MW);
    SyncConst4 sc;
    sc.kmaskhi = a.k > 16 ? (a.k == 32 ? 0xFFFFFFFFu : ((1u << (2 * a.k - 32)) - 1u)) : 0u;
    sc.kmasklo = a.k >= 16 ? 0xFFFFFFFFu : ((1u << (2 * a.k)) - 1u);
    // an offset no window has (>= w) never matches: bytes of d are at most B = w - 1 < 0x7F
    const u32 s = (u32)a.sync_s < (u32)a.w ? (u32)a.sync_s : 0x7Fu, e = (u32)a.sync_e < (u32)a.w ? (u32)a.sync_e : 0x7Fu;
    sc.tF0 = s * 0x01010101u; sc.tF1 = e * 0x01010101u;
    sc.tR0 = (s == 0x7Fu ? 0x7Fu : (u32)a.w - 1u - s) * 0x01010101u;
    sc.tR1 = (e == 0x7Fu ? 0x7Fu : (u32)a.w - 1u - e) * 0x01010101u;
    tile_loop4<G>(a, wsm, t,
                  [&](int b, long long tile, int have_pre) { return compute_sync4<B, NB, LW, CANON, MW>(a, wsm, c, sc, b, tile, t, have_pre); },
                  [&](int b, long long tile, u32 n_rec, u64 base) { emit_sync4<B, NB, CANON>(a, wsm, c, sc, b, tile, t, n_rec, base); });
}

}  // namespace blg
