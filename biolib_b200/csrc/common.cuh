// common.cuh -- device helpers shared by the kernels of the streaming k-mer path (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "nt4_swar.h"

namespace blg {

typedef unsigned long long u64;
typedef unsigned int u32;
typedef unsigned char u8;

// ---------------------------------------------------------------------------------------------
// MurmurHash3_x64_128 specialised to an 8-byte key, low word only: what hash::hash64::hash<uint64_t>
// returns (reference include/hash.hpp:55-59 -> :16-21 -> bundled/MurmurHash3.cpp:263-341, tail case 8
// at :314-322, finalisation :328-337, fmix64 :81-90). For len == 8 the second lane h2 is still the
// seed when finalisation starts, so s8 = (uint64)(uint32)seed ^ 8 folds both `^= len` steps.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ u64 rotl64(u64 x, int r) { return (x << r) | (x >> (64 - r)); }

#ifdef BLG_HASH_PLAIN
// The arithmetic as written in the reference, lowered by nvcc (development A/B builds only): 48 instructions per hash, 25 of them
// on the ALU pipe.
__device__ __forceinline__ u64 fmix64(u64 h)
{
    h ^= h >> 33; h *= 0xff51afd7ed558ccdULL;
    h ^= h >> 33; h *= 0xc4ceb9fe1a85ec53ULL;
    return h ^ (h >> 33);
}
__device__ __forceinline__ u64 murmur3_key8_s32(u64 key, u32 s8)
{
    u64 k1 = key * 0x87c37b91114253d5ULL;
    k1 = rotl64(k1, 31);
    k1 *= 0x4cf5ad432745937fULL;
    u64 h1 = k1 ^ (u64)s8;      // (seed ^ k1) ^ len
    u64 h2 = (u64)s8;           // seed ^ len
    h1 += h2; h2 += h1;
    h1 = fmix64(h1); h2 = fmix64(h2);
    return h1 + h2;
}
#else
// The same arithmetic with the lowering chosen by hand, because the fused kernels are bound by the ALU pipe (LOP3 / SHF / IADD3)
// while the FMA pipe (IMAD) has room: 39 instructions per hash, 18 of them on the ALU pipe in the fused kernel.
//   x * c (low 64 bits) = IMAD.WIDE of the low words + two IMADs that accumulate the cross products into the high word. nvcc's
//   own lowering computes the cross products beside the wide product and joins them with an IADD3 (shorter dependency chain,
//   one more instruction, on the ALU pipe);
//   64-bit + 32-bit written as mad.wide.u32 a * 1 + c: ptxas emits IADD3 + IMAD.X (the carry into the high word goes to the FMA
//   pipe) where the plain C addition gave IADD3 + IADD3.X;
//   h ^= h >> 33 only touches the low word: lo ^= hi >> 1.
// Measured on the fused kernel (profiles/r02_zn_hash_lowering_ab.txt): 209 -> 219 Gbases/s at k=31 m=19, 181 -> 198 on 150-bp reads.
__device__ __forceinline__ void mul64c(u32& xl, u32& xh, u64 c)
{
    u32 pl, ph;
    asm("{\n\t.reg .u64 p;\n\tmul.wide.u32 p, %2, %4;\n\tmov.b64 {%0, %1}, p;\n\tmad.lo.u32 %1, %3, %4, %1;\n\tmad.lo.u32 %1, %2, %5, %1;\n\t}"
        : "=&r"(pl), "=&r"(ph) : "r"(xl), "r"(xh), "r"((u32)c), "r"((u32)(c >> 32)));
    xl = pl; xh = ph;
}
__device__ __forceinline__ u64 add32w(u32 a, u64 c)
{
    u64 r;
    asm("mad.wide.u32 %0, %1, 1, %2;" : "=l"(r) : "r"(a), "l"(c));
    return r;
}
// (h >> 1 as IMAD.HI by 2^31 on the FMA pipe instead of SHF on the ALU pipe: measured again with this lowering, slower at 2, 4 and 6
// of the six sites -- 0.454 -> 0.459 / 0.472 / 0.485 ms per 100 Mbp, profiles/r02_zs_hash_imad_hi_ab.txt.)
__device__ __forceinline__ void fmix_step(u32& l, u32& h, u64 c) { l ^= h >> 1; mul64c(l, h, c); }

// s8 = (uint32)seed ^ 8: the seed is 32 bits wide (hash.hpp:50), so the high words of both `^= s8` / `+= s8` vanish
__device__ __forceinline__ u64 murmur3_key8_s32(u64 key, u32 s8)
{
    u32 l = (u32)key, h = (u32)(key >> 32);
    mul64c(l, h, 0x87c37b91114253d5ULL);
    { const u32 nl = __funnelshift_l(h, l, 31), nh = __funnelshift_l(l, h, 31); l = nl; h = nh; }   // rotl64(k1, 31)
    mul64c(l, h, 0x4cf5ad432745937fULL);
    const u64 h1 = add32w(s8, ((u64)h << 32) | (l ^ s8));      // h1 = (k1 ^ s8) + h2, h2 = s8
    const u64 h2 = add32w(s8, h1);                             // h2 += h1
    u32 l1 = (u32)h1, u1 = (u32)(h1 >> 32), l2 = (u32)h2, u2 = (u32)(h2 >> 32);
    fmix_step(l1, u1, 0xff51afd7ed558ccdULL); fmix_step(l2, u2, 0xff51afd7ed558ccdULL);
    fmix_step(l1, u1, 0xc4ceb9fe1a85ec53ULL); fmix_step(l2, u2, 0xc4ceb9fe1a85ec53ULL);
    l1 ^= u1 >> 1; l2 ^= u2 >> 1;
    return add32w(l2, ((u64)u1 << 32) | l1) + ((u64)u2 << 32);
}
#endif

__device__ __forceinline__ u64 murmur3_key8(u64 key, u64 s8) { return murmur3_key8_s32(key, (u32)s8); }

// ---------------------------------------------------------------------------------------------
// One-dimensional bulk async copy global -> shared (cp.async.bulk, the TMA unit without a tensor map) with completion on an
// mbarrier: the staging the fused kernels and kmer2_kernel use for the next tile's ASCII bytes.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* b, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, u32 bytes, unsigned long long* b)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* b, u32 phase)
{
    asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}"
                 ::"r"(smem_u32(b)), "r"(phase) : "memory");
}
// generic-proxy accesses to shared memory before this point are ordered before async-proxy (bulk copy) writes after it
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---------------------------------------------------------------------------------------------
// 2-bit streams. Bases are packed 16 per 32-bit word, base q of a word at bits [2q, 2q+2)
// (little-endian stream: earlier base = lower bits). With that geometry
//   reverse-complement word of bases [p, p+len)  = ~stream & mask            (free)
//   forward word (first base most significant)   = pair-reverse of the stream (one BREV + swap)
// which are the `rev` / `fwd` words of reference include/kmer_view.hpp:194-195 and
// include/minimizer_view.hpp:236-237.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ u64 pairrev64(u64 x)
{
    x = __brevll(x);
    return ((x >> 1) & 0x5555555555555555ULL) | ((x & 0x5555555555555555ULL) << 1);
}

// tile staging word: x = 16 codes, y = 16 flag pairs (bit0 = not ACGTU, bit1 = a sequence starts here)
__device__ __forceinline__ u64 stream64_codes(const uint2* cf, int base_idx)
{
    const int wi = base_idx >> 4, sh = (base_idx & 15) * 2;
    const u32 a = cf[wi].x, b = cf[wi + 1].x, c = cf[wi + 2].x;
    return ((u64)__funnelshift_r(b, c, sh) << 32) | __funnelshift_r(a, b, sh);
}
__device__ __forceinline__ u64 stream64_flags(const uint2* cf, int base_idx)
{
    const int wi = base_idx >> 4, sh = (base_idx & 15) * 2;
    const u32 a = cf[wi].y, b = cf[wi + 1].y, c = cf[wi + 2].y;
    return ((u64)__funnelshift_r(b, c, sh) << 32) | __funnelshift_r(a, b, sh);
}

__device__ __forceinline__ u64 lowmask64(int bits) { return bits >= 64 ? ~0ULL : ((1ULL << bits) - 1); }

// ---------------------------------------------------------------------------------------------
// Staging: ASCII bases [A0, A0 + 16*nw) -> cf[0..nw) (+2 all-bad pad words). Positions outside
// [buf_lo, buf_hi) read as byte 0, which the table maps to "break". 16-byte vector loads when the
// buffer's position 0 is 16-byte aligned (vec_ok), byte loads otherwise and at the edges.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void stage_tile(uint2* cf, int nw, const u8* __restrict__ bases, long long buf_lo,
                                           long long buf_hi, long long A0, int vec_ok, int tid, int nthreads)
{
    for (int wi = tid; wi < nw + 2; wi += nthreads) {
        u32 codes = 0, flags = 0x55555555u;
        if (wi < nw) {
            const long long P = A0 + 16LL * wi;
            u32 w0, w1, w2, w3;
            if (vec_ok && P >= buf_lo && P + 16 <= buf_hi) {
                const uint4 v = __ldg(reinterpret_cast<const uint4*>(bases + (P - buf_lo)));
                w0 = v.x; w1 = v.y; w2 = v.z; w3 = v.w;
            } else {
                u32 ww[4] = {0, 0, 0, 0};
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    const long long pos = P + q;
                    if (pos >= buf_lo && pos < buf_hi) ww[q >> 2] |= (u32)bases[pos - buf_lo] << (8 * (q & 3));
                }
                w0 = ww[0]; w1 = ww[1]; w2 = ww[2]; w3 = ww[3];
            }
            blg_pack16(w0, w1, w2, w3, &codes, &flags);
        }
        cf[wi] = make_uint2(codes, flags);
    }
}

// Mark the sequence starts that fall inside the staged range (interior offsets only): one warp.
__device__ __forceinline__ void mark_sequence_starts(uint2* cf, int nw, const u64* __restrict__ offs, long long n_seq,
                                                     long long A0, int lane)
{
    // first q in [1, n_seq) with offs[q] >= A0
    long long lo = 1, hi = n_seq;
    const long long lo_pos = A0 < 0 ? 0 : A0, end_pos = A0 + 16LL * nw;
    while (lo < hi) {
        const long long mid = (lo + hi) >> 1;
        if ((long long)offs[mid] < lo_pos) lo = mid + 1; else hi = mid;
    }
    for (long long q = lo + lane; q < n_seq; q += 32) {
        const long long o = (long long)offs[q];
        if (o >= end_pos) break;
        const int rel = (int)(o - A0);
        atomicOr(&cf[rel >> 4].y, 2u << (2 * (rel & 15)));
    }
}

// First q in [1, n_seq) with offs[q] >= pos (n_seq if none), found by the whole warp: one probe of 32 consecutive
// entries around the position-proportional guess (exact for equal-length reads), then 32-ary narrowing (5 rounds for
// 10^8 sequences instead of 27 dependent loads per lane).
__device__ __forceinline__ long long first_offset_at_or_after(const u64* __restrict__ offs, long long n_seq, long long pos,
                                                              long long total, int t)
{
    long long lo = 1, hi = n_seq;                      // the answer lies in [lo, hi]
    if (hi <= lo) return hi;
    {
        long long g = total > 0 ? (long long)((double)pos / (double)total * (double)n_seq) - 15 : 1;
        g = g < 1 ? 1 : (g > n_seq - 1 ? n_seq - 1 : g);
        const long long q = g + t;
        const bool ge = q >= n_seq || (long long)offs[q] >= pos;
        const unsigned m = __ballot_sync(0xffffffffu, ge);
        if (m == 0) lo = g + 32 < hi ? g + 32 : hi;               // everything probed is below pos
        else {
            const int f = __ffs((int)m) - 1;
            if (f > 0 || g == 1) return (g + f) < n_seq ? g + f : n_seq;   // the entry before it was probed and is below
            hi = g;                                               // the first probe already is >= pos
        }
    }
    while (hi - lo > 0) {
        const long long span = hi - lo;                // candidates lo .. hi-1 (hi itself is known to satisfy or is n_seq)
        const long long step = (span + 31) / 32;
        const long long q = lo + (long long)t * step;
        const bool ge = q >= hi || (long long)offs[q] >= pos;
        const unsigned m = __ballot_sync(0xffffffffu, ge);
        if (m == 0) { lo = lo + 31 * step + 1 < hi ? lo + 31 * step + 1 : hi; continue; }   // all probed are below
        const int f = __ffs((int)m) - 1;
        const long long qf = lo + (long long)f * step;             // first probed entry that is >= pos
        if (step == 1) return qf < hi ? qf : hi;
        hi = qf < hi ? qf : hi;
        lo = f > 0 ? qf - step + 1 : lo;
    }
    return hi;
}

// ---------------------------------------------------------------------------------------------
// Cross-tile ordered compaction: single-pass chained scan with decoupled look-back.
// status word = flag (bits 63..62: 1 = tile aggregate, 2 = inclusive prefix) | 62-bit count.
// Tile ids come from an atomic ticket, so every lower tile is held by a CTA that is already
// running: the spin below cannot deadlock. Called by one full warp; returns the exclusive prefix.
// ---------------------------------------------------------------------------------------------
static constexpr u64 ST_AGG = 1ULL << 62, ST_INCL = 2ULL << 62, ST_VAL = (1ULL << 62) - 1;

__device__ __forceinline__ u64 ld_status(const u64* p)
{
    u64 v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_status(u64* p, u64 v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ u64 lookback_exclusive(u64* status, long long tile, u64 aggregate, int lane)
{
    const unsigned full = 0xffffffffu;
    if (tile == 0) {
        if (lane == 0) st_status(&status[0], ST_INCL | aggregate);
        return 0;
    }
    if (lane == 0) st_status(&status[tile], ST_AGG | aggregate);
    u64 excl = 0;
    long long look = tile - 1;
    for (;;) {
        const long long idx = look - lane;
        u64 st = idx >= 0 ? ld_status(&status[idx]) : ST_INCL;
        while (__any_sync(full, (st >> 62) == 0)) {
            if ((st >> 62) == 0) st = ld_status(&status[idx]);
        }
        const unsigned incl = __ballot_sync(full, (st >> 62) == 2);
        const int first_incl = incl ? (__ffs(incl) - 1) : 32;
        u64 c = lane <= first_incl ? (st & ST_VAL) : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(full, c, o);
        excl += c;
        if (incl) break;
        look -= 32;
    }
    if (lane == 0) st_status(&status[tile], ST_INCL | (excl + aggregate));
    return excl;
}

// Split, wide variant used by sk2_kernel: a scan unit first PUBLISHES its aggregate (one store, nobody waits), and
// RESOLVES its exclusive prefix later -- half a tile of other work later -- when every predecessor's aggregate has
// long been out, so the walk below practically never polls. 128 predecessors per round (four per lane, nearest first).
__device__ __forceinline__ void scan_publish(u64* status, long long unit, u64 aggregate)
{
    // atomicMax: the inclusive word (flag 2) written by whichever warp resolves this unit can never be overwritten
    // by a late aggregate (flag 1)
    atomicMax(&status[unit], (unit == 0 ? ST_INCL : ST_AGG) | aggregate);
}

__device__ __forceinline__ u64 scan_resolve(u64* status, long long unit, u64 aggregate, int lane)
{
    const unsigned full = 0xffffffffu;
    if (unit == 0) return 0;
    u64 excl = 0;
    long long look = unit - 1;
    for (;;) {
        u64 sum = 0;
        int found = 0, pending = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const long long idx = look - (4 * lane + j);
            const u64 st = idx >= 0 ? ld_status(&status[idx]) : ST_INCL;
            const u32 fl = (u32)(st >> 62);
            if (!found && !pending) {
                if (fl == 0) pending = 1;
                else { sum += st & ST_VAL; found = (fl == 2); }
            }
        }
        const unsigned stop = __ballot_sync(full, found | pending);
        const int first = stop ? (__ffs(stop) - 1) : 32;
        const int first_pending = __shfl_sync(full, pending, first & 31) && first < 32;
        if (first_pending) { __nanosleep(40); continue; }   // the nearest undecided word is still empty: poll again
        u64 c = lane <= first ? sum : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(full, c, o);
        excl += c;
        if (first < 32) break;
        look -= 128;
    }
    if (lane == 0) st_status(&status[unit], ST_INCL | (excl + aggregate));
    return excl;
}

}  // namespace blg
