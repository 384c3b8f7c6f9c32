// rank_kernel.cu -- minimizer_context_t::id on the device (only when BLG_F_EMIT_ID is requested).
//
// SURVEY D9 (reference include/minimizer_view.hpp:257,263: `mmer_count` is stamped into each m-mer as it is pushed
// and incremented per pushed m-mer): id of a minimizer = number of VALID m-mer starts before it inside its sequence.
// That is a rank query, so the pass builds a rank directory over the per-position "m-mer starts here and is valid"
// bit vector and answers one query per record:
//   valid_words   one bit per position, 64 positions per word                      (n/8 bytes)
//   word_rel      exclusive popcount prefix of each word inside its 65536-position superblock (u32 per word)
//   super_prefix  exclusive popcount prefix of each superblock                      (u64 per 65536 positions)
//   rank(j) = super_prefix[j >> 16] + word_rel[j >> 6] + popc(valid_words[j >> 6] & low(j & 63))
//   id = rank(mm_start) - rank(start of the sequence that holds mm_start)
// One extra read of the bases (1 B/base) plus ~0.2 B/base of directory traffic; none of it is counted as
// algorithmic bytes of the path.
#include "common.cuh"
#include "kernels.h"

namespace blg {

namespace {

constexpr int RK_THREADS = 1024;          // words per superblock = threads per CTA
constexpr int RK_SUPER_SHIFT = 16;        // 65536 positions

__device__ __forceinline__ u32 compress_even_bits(u32 x)
{
    x &= 0x55555555u;
    x = (x | (x >> 1)) & 0x33333333u;
    x = (x | (x >> 2)) & 0x0F0F0F0Fu;
    x = (x | (x >> 4)) & 0x00FF00FFu;
    x = (x | (x >> 8)) & 0x0000FFFFu;
    return x;
}

// bad-byte bits of the 16 bases starting at position P (positions outside [0, n) are bad)
__device__ __forceinline__ u32 bad16(const u8* __restrict__ bases, long long n, long long P, int vec_ok)
{
    u32 w0, w1, w2, w3;
    if (vec_ok && P >= 0 && P + 16 <= n) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(bases + P));
        w0 = v.x; w1 = v.y; w2 = v.z; w3 = v.w;
    } else {
        u32 ww[4] = {0, 0, 0, 0};
#pragma unroll
        for (int q = 0; q < 16; ++q) {
            const long long pos = P + q;
            if (pos >= 0 && pos < n) ww[q >> 2] |= (u32)bases[pos] << (8 * (q & 3));
        }
        w0 = ww[0]; w1 = ww[1]; w2 = ww[2]; w3 = ww[3];
    }
    u32 codes, flags;
    blg_pack16(w0, w1, w2, w3, &codes, &flags);
    return compress_even_bits(flags);
}

// one thread = one word of 64 m-mer starts: valid iff none of the m bytes from the start is bad
__global__ void __launch_bounds__(256) valid_words_kernel(const u8* __restrict__ bases, long long n, int m, int vec_ok,
                                                         u64* __restrict__ valid_words, long long n_words)
{
    typedef unsigned __int128 u128;
    for (long long wq = blockIdx.x * (long long)blockDim.x + threadIdx.x; wq < n_words; wq += (long long)gridDim.x * blockDim.x) {
        const long long P = wq * 64;
        u128 bad = 0;
#pragma unroll
        for (int g = 0; g < 6; ++g) bad |= (u128)bad16(bases, n, P + 16 * g, vec_ok) << (16 * g);
        // invalid(j) = OR over d < m of bad(j + d): doubling OR of right shifts
        u128 acc = bad;
        int covered = 1;
        while (covered < m) {
            const int step = covered < m - covered ? covered : m - covered;
            acc |= acc >> step;
            covered += step;
        }
        valid_words[wq] = ~(u64)acc;
    }
}

// m-mers may not run across a sequence end: clear the m-1 starts before every interior offset
__global__ void clear_boundaries_kernel(const u64* __restrict__ offs, long long n_seq, int m, u64* valid_words)
{
    for (long long q = 1 + blockIdx.x * (long long)blockDim.x + threadIdx.x; q < n_seq; q += (long long)gridDim.x * blockDim.x) {
        const long long o = (long long)offs[q];
        long long lo = o - (m - 1);
        if (lo < 0) lo = 0;
        if (lo >= o) continue;
        // bits [lo, o): at most two words since m - 1 <= 31
        const long long w0 = lo >> 6, w1 = (o - 1) >> 6;
        const u64 from = ~0ULL << (lo & 63);
        const u64 upto = ((o & 63) == 0) ? ~0ULL : ((1ULL << (o & 63)) - 1);
        if (w0 == w1) atomicAnd(&valid_words[w0], ~(from & upto));
        else {
            atomicAnd(&valid_words[w0], ~from);
            atomicAnd(&valid_words[w1], ~upto);
        }
    }
}

// one CTA = one superblock of 1024 words: exclusive popcount prefix per word + the superblock total
__global__ void __launch_bounds__(RK_THREADS) word_prefix_kernel(const u64* __restrict__ valid_words, long long n_words,
                                                                u32* __restrict__ word_rel, u64* __restrict__ super_total)
{
    __shared__ u32 wsum[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long wq = (long long)blockIdx.x * RK_THREADS + tid;
    const u32 c = wq < n_words ? (u32)__popcll(valid_words[wq]) : 0u;
    u32 incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const u32 v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        u32 v = wsum[lane], s = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const u32 t = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += t;
        }
        wsum[lane] = s - v;
        if (lane == 31) super_total[blockIdx.x] = s;
    }
    __syncthreads();
    if (wq < n_words) word_rel[wq] = wsum[warp] + incl - c;
}

// exclusive scan of the superblock totals, in place; one CTA walks the array in chunks of 1024
__global__ void __launch_bounds__(RK_THREADS) super_scan_kernel(u64* super, long long n_super)
{
    __shared__ u64 wsum[32];
    __shared__ u64 carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (long long base = 0; base < n_super; base += RK_THREADS) {
        const long long i = base + tid;
        const u64 c = i < n_super ? super[i] : 0ULL;
        u64 incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const u64 v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            u64 v = wsum[lane], s = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const u64 t = __shfl_up_sync(0xffffffffu, s, o);
                if (lane >= o) s += t;
            }
            wsum[lane] = s - v;
        }
        __syncthreads();
        const u64 excl = carry + wsum[warp] + incl - c;
        if (i < n_super) super[i] = excl;
        __syncthreads();
        if (tid == RK_THREADS - 1) carry = excl + c;
        __syncthreads();
    }
}

__device__ __forceinline__ u64 rank_at(const u64* __restrict__ valid_words, const u32* __restrict__ word_rel,
                                       const u64* __restrict__ super_prefix, u64 j)
{
    const u64 wq = j >> 6;
    const u64 low = (j & 63) ? (valid_words[wq] & ((1ULL << (j & 63)) - 1)) : 0ULL;
    return super_prefix[j >> RK_SUPER_SHIFT] + word_rel[wq] + (u64)__popcll(low);
}

__global__ void __launch_bounds__(256) record_id_kernel(const u64* __restrict__ mm_start, u64 n_records,
                                                       const u64* __restrict__ offs, long long n_seq,
                                                       const u64* __restrict__ valid_words, const u32* __restrict__ word_rel,
                                                       const u64* __restrict__ super_prefix, u64* __restrict__ id)
{
    for (u64 r = blockIdx.x * (u64)blockDim.x + threadIdx.x; r < n_records; r += (u64)gridDim.x * blockDim.x) {
        const u64 j = mm_start[r];
        u64 base = 0;
        if (offs != nullptr && n_seq > 1) {
            long long lo = 0, hi = n_seq;  // last q with offs[q] <= j
            while (hi - lo > 1) {
                const long long mid = (lo + hi) >> 1;
                if (offs[mid] <= j) lo = mid; else hi = mid;
            }
            const u64 s0 = offs[lo];
            if (s0) base = rank_at(valid_words, word_rel, super_prefix, s0);
        }
        id[r] = rank_at(valid_words, word_rel, super_prefix, j) - base;
    }
}

}  // namespace

size_t rank_scratch_bytes(long long n_bases)
{
    const long long n_words = (n_bases + 63) / 64 + 1;
    const long long n_super = (n_words + RK_THREADS - 1) / RK_THREADS;
    return (size_t)n_words * 8 + (size_t)n_words * 4 + (size_t)n_super * 8 + 64;
}

cudaError_t record_ids_launch(const unsigned char* bases, long long n_bases, const unsigned long long* offs, long long n_seq,
                              int m, const unsigned long long* mm_start, unsigned long long n_records, void* scratch,
                              unsigned long long* id, cudaStream_t stream, int* launches)
{
    if (n_records == 0) return cudaSuccess;
    const long long n_words = (n_bases + 63) / 64 + 1;   // +1: rank(n_bases) of a sequence start at the very end stays in range
    const long long n_super = (n_words + RK_THREADS - 1) / RK_THREADS;
    u64* valid_words = static_cast<u64*>(scratch);
    u64* super = valid_words + n_words;
    u32* word_rel = reinterpret_cast<u32*>(super + n_super);
    const int vec_ok = (reinterpret_cast<uintptr_t>(bases) & 15u) == 0;
    long long blocks = (n_words + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    valid_words_kernel<<<(unsigned)blocks, 256, 0, stream>>>(bases, n_bases, m, vec_ok, valid_words, n_words);
    int nl = 1;
    if (offs != nullptr && n_seq > 1 && m > 1) {
        long long b2 = (n_seq + 255) / 256;
        if (b2 > 148 * 8) b2 = 148 * 8;
        clear_boundaries_kernel<<<(unsigned)b2, 256, 0, stream>>>(offs, n_seq, m, valid_words);
        ++nl;
    }
    word_prefix_kernel<<<(unsigned)n_super, RK_THREADS, 0, stream>>>(valid_words, n_words, word_rel, super);
    super_scan_kernel<<<1, RK_THREADS, 0, stream>>>(super, n_super);
    unsigned long long b3 = (n_records + 255) / 256;
    if (b3 > 148 * 8) b3 = 148 * 8;
    record_id_kernel<<<(unsigned)b3, 256, 0, stream>>>(mm_start, n_records, offs, n_seq, valid_words, word_rel, super, id);
    nl += 3;
    if (launches) *launches += nl;
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// hash::minimizer_position_extractor, batched (reference include/kmer_view.hpp:266-283, ctor src/kmer_view.cpp:7-15):
// one thread per k-mer value (k <= 32). The k-m+1 m-mers are visited from the right (km >>= 2), hashed with seed 0,
// and a later (further left) m-mer replaces the minimum on '>=' -- smallest hash, leftmost on ties. out = k - m - minpos.
// Algorithmic bytes: 8 in + 1 out per k-mer; w hashes per k-mer make it ALU-bound like the fused kernels.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) minpos_kernel(const u64* __restrict__ values, size_t n, int k, int m, unsigned char* __restrict__ out)
{
    const u64 mask = (2 * m != 64) ? ((1ULL << (2 * m)) - 1ULL) : ~0ULL;
    const int w = k - m + 1;
    for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (size_t)gridDim.x * blockDim.x) {
        u64 km = values[q];
        u64 mval = ~0ULL;
        int minpos = 0;
        for (int i = 0; i < w; ++i) {
            const u64 val = murmur3_key8(km & mask, 8ULL);     // seed 0: s8 = 0 ^ 8
            if (mval >= val) { mval = val; minpos = i; }
            km >>= 2;
        }
        out[q] = (unsigned char)(k - m - minpos);
    }
}

cudaError_t minpos_launch(const unsigned long long* values, size_t n, int k, int m, unsigned char* out, int sm_count, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    size_t blocks = (n + 255) / 256;
    const size_t cap = (size_t)sm_count * 8;
    if (blocks > cap) blocks = cap;
    minpos_kernel<<<(unsigned)blocks, 256, 0, stream>>>(values, n, k, m, out);
    return cudaGetLastError();
}

}  // namespace blg
