// payload_kernel.cu -- what super-k-mers are for: every record's bases as 2-bit codes, grouped by
// minimizer bucket, ready for partitioned k-mer counting / de Bruijn graph construction.
//
//   bucket(record) = hash64(minimizer value, seed) % n_buckets      (the hash that chose the minimizer: include/hash.hpp:55-59)
//   records are reordered by bucket, STABLY (inside a bucket they keep the stream order), and the bases
//   [sk_start, sk_start + size + k - 1) of every record are appended, in that order, to ONE 2-bit packed stream laid out like
//   the reference's bit::packed::vector<uint64_t>(2) (include/packed_vector.hpp:134-160,311-318): element i sits in word
//   i / 32 at bits [62 - 2 (i % 32), 64 - 2 (i % 32)), i.e. first base most significant, with A=0 C=1 G=2 T=3
//   (constants::seq_nt4_table).
//
// Passes (all record-parallel, the records being device-resident output of the fused kernel):
//   count    tiles of 2048 records: per-tile bucket histogram in shared memory -> matrix M[bucket][tile]
//   scan     exclusive prefix sum of M in bucket-major order: where every (bucket, tile) group starts
//   scatter  same tiles; the rank of a record inside its group comes from ballots (lanes of a warp with the same bucket) and a
//            packed per-warp counter per bucket, so the order is the stream order; writes the reordered record arrays and every
//            record's payload length
//   scan     exclusive prefix sum of the lengths: the payload start of every record in the packed stream
//   pack     one thread per record: its ASCII bases as aligned 8-byte words, eight bases -> sixteen stream bits at a time, ORed into
//            the stream (a record shares its first and last word with its neighbours: atomicOr there, plain stores in between)
#include "common.cuh"
#include "kernels.h"

namespace blg {

namespace {

constexpr int PL_THREADS = 256;
constexpr int PL_CHUNKS = 8;
constexpr int PL_TILE = PL_THREADS * PL_CHUNKS;     // records per tile

__device__ __forceinline__ u32 bucket_of(u64 value, u64 s8, u32 n_buckets) { return (u32)(murmur3_key8(value, s8) % n_buckets); }

__global__ void __launch_bounds__(PL_THREADS) pl_count(const u64* __restrict__ value, u64 n, u64 s8, u32 n_buckets, u64 n_tiles, u64* __restrict__ M)
{
    extern __shared__ u32 hist[];
    for (u32 b = threadIdx.x; b < n_buckets; b += PL_THREADS) hist[b] = 0;
    __syncthreads();
    const u64 base = (u64)blockIdx.x * PL_TILE;
#pragma unroll
    for (int c = 0; c < PL_CHUNKS; ++c) {
        const u64 r = base + (u64)c * PL_THREADS + threadIdx.x;
        if (r < n) atomicAdd(&hist[bucket_of(__ldg(value + r), s8, n_buckets)], 1u);
    }
    __syncthreads();
    for (u32 b = threadIdx.x; b < n_buckets; b += PL_THREADS) M[(u64)b * n_tiles + blockIdx.x] = hist[b];
}

// ---- exclusive scan of u64 arrays: per-block sums, one-CTA scan of the sums, apply ------------------------------------
constexpr int SC_THREADS = 256, SC_PER = 8, SC_BLOCK = SC_THREADS * SC_PER;

__global__ void __launch_bounds__(SC_THREADS) sc_sums(const u64* __restrict__ a, u64 n, u64* __restrict__ sums)
{
    __shared__ u64 s[SC_THREADS / 32];
    const u64 base = (u64)blockIdx.x * SC_BLOCK + (u64)threadIdx.x * SC_PER;
    u64 t = 0;
#pragma unroll
    for (int i = 0; i < SC_PER; ++i) if (base + i < n) t += a[base + i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) { u64 x = 0; for (int w = 0; w < SC_THREADS / 32; ++w) x += s[w]; sums[blockIdx.x] = x; }
}

__global__ void __launch_bounds__(1024) sc_scan_sums(u64* __restrict__ sums, u64 n_blocks, u64* __restrict__ total)
{
    __shared__ u64 s[1024];
    const int tid = threadIdx.x;
    const u64 per = (n_blocks + 1023) / 1024;
    const u64 lo = min(n_blocks, (u64)tid * per), hi = min(n_blocks, lo + per);
    u64 t = 0;
    for (u64 i = lo; i < hi; ++i) t += sums[i];
    s[tid] = t;
    __syncthreads();
    if (tid == 0) { u64 r = 0; for (int i = 0; i < 1024; ++i) { const u64 x = s[i]; s[i] = r; r += x; } *total = r; }
    __syncthreads();
    u64 r = s[tid];
    for (u64 i = lo; i < hi; ++i) { const u64 x = sums[i]; sums[i] = r; r += x; }
}

__global__ void __launch_bounds__(SC_THREADS) sc_apply(u64* __restrict__ a, u64 n, const u64* __restrict__ sums)
{
    __shared__ u64 s[SC_THREADS];
    const u64 base = (u64)blockIdx.x * SC_BLOCK + (u64)threadIdx.x * SC_PER;
    u64 v[SC_PER], t = 0;
#pragma unroll
    for (int i = 0; i < SC_PER; ++i) { v[i] = base + i < n ? a[base + i] : 0; t += v[i]; }
    s[threadIdx.x] = t;
    __syncthreads();
    // exclusive prefix over the threads of the block (256 entries: a short serial walk per thread over shared memory)
    u64 r = sums[blockIdx.x];
    for (int j = 0; j < (int)threadIdx.x; ++j) r += s[j];
#pragma unroll
    for (int i = 0; i < SC_PER; ++i) { if (base + i < n) a[base + i] = r; r += v[i]; }
}

__device__ __forceinline__ u32 sum_bytes8(unsigned long long x)
{
    unsigned long long t = (x & 0x00FF00FF00FF00FFULL) + ((x >> 8) & 0x00FF00FF00FF00FFULL);
    t = (t & 0x0000FFFF0000FFFFULL) + ((t >> 16) & 0x0000FFFF0000FFFFULL);
    return (u32)(t & 0xFFFFFFFFULL) + (u32)(t >> 32);
}

// ---- scatter: stable rank inside (bucket, tile) ---------------------------------------------------------------------------
// A chunk is 256 consecutive records, 32 per warp. Lanes of a warp with the same bucket are found with ballots over the bucket's
// bits; the leader of every such group adds its size into byte `warp` of a packed 64-bit counter of the bucket (a group has at most
// 32 members, a byte holds it): after one barrier the bytes below `warp` are the records of earlier warps of the chunk, all eight
// the chunk's total. Three barriers per chunk instead of one per warp.
__global__ void __launch_bounds__(PL_THREADS) pl_scatter(const u64* __restrict__ value, const u64* __restrict__ mm_start, const u8* __restrict__ size,
                                                         const u8* __restrict__ mm_pos, u64 n, u64 s8, u32 n_buckets, int bucket_bits, u64 n_tiles, int k,
                                                         const u64* __restrict__ M, u64* __restrict__ o_value, u64* __restrict__ o_mm,
                                                         u8* __restrict__ o_size, u8* __restrict__ o_pos, u64* __restrict__ o_len)
{
    extern __shared__ u64 pl_smem[];
    unsigned long long* pk = reinterpret_cast<unsigned long long*>(pl_smem);   // [n_buckets] this chunk: records per warp, one byte each
    u32* cnt = reinterpret_cast<u32*>(pl_smem + n_buckets);                     // [n_buckets] records of this tile already placed
    for (u32 b = threadIdx.x; b < n_buckets; b += PL_THREADS) { cnt[b] = 0; pk[b] = 0; }
    __syncthreads();
    const u64 base = (u64)blockIdx.x * PL_TILE;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int c = 0; c < PL_CHUNKS; ++c) {
        const u64 r = base + (u64)c * PL_THREADS + threadIdx.x;
        const bool have = r < n;
        u64 v = 0;
        u32 b = 0;
        if (have) { v = __ldg(value + r); b = bucket_of(v, s8, n_buckets); }
        u32 peers = __ballot_sync(0xffffffffu, have);
        for (int bit = 0; bit < bucket_bits; ++bit) {
            const u32 x = __ballot_sync(0xffffffffu, (b >> bit) & 1u);
            peers &= ((b >> bit) & 1u) ? x : ~x;
        }
        const u32 rank_in_warp = __popc(peers & ((1u << lane) - 1u));
        const bool leader = have && rank_in_warp == 0;
        if (leader) atomicAdd(&pk[b], (unsigned long long)__popc(peers) << (8 * warp));
        __syncthreads();
        u32 first = 0, chunk_total = 0;
        bool closes = false;
        if (have) {
            const unsigned long long packed = pk[b];
            const unsigned long long below = packed & ((1ULL << (8 * warp)) - 1ULL);
            first = cnt[b] + sum_bytes8(below);
            chunk_total = sum_bytes8(packed);
            closes = leader && below == 0;         // the first warp of the chunk that holds this bucket closes it after the barrier
        }
        __syncthreads();
        if (closes) { cnt[b] += chunk_total; pk[b] = 0; }
        if (have) {
            const u64 dst = M[(u64)b * n_tiles + blockIdx.x] + first + rank_in_warp;
            const u8 sz = size[r];
            o_value[dst] = v;
            o_mm[dst] = mm_start[r];
            o_size[dst] = sz;
            o_pos[dst] = mm_pos[r];
            o_len[dst] = (u64)sz + (u64)(k - 1);
        }
        __syncthreads();
    }
}

// bucket_offsets[b] = first record of bucket b = M[b][0] after the scan; [n_buckets] = n
__global__ void pl_bucket_offsets(const u64* __restrict__ M, u64 n_tiles, u32 n_buckets, u64 n, u64* __restrict__ out)
{
    const u32 b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < n_buckets) out[b] = M[(u64)b * n_tiles];
    if (b == n_buckets) out[b] = n;
}

// ---- pack: one thread per (reordered) record ---------------------------------------------------------------------------------
// The record's bases are read as aligned 8-byte words (the granule that holds a valid byte is always readable), eight bases become
// sixteen stream bits at a time (SWAR codes, then the pairs reversed so that the first base is on top), and are shifted into the
// 64-bit words of the stream.
__device__ __forceinline__ u32 codes8_msb_first(u64 w)
{
    const u32 c = blg_codes4((u32)w) | (blg_codes4((u32)(w >> 32)) << 8);      // base i at bits [2i, 2i+2)
    const u32 r = __brev(c) >> 16;                                               // base i at bits [14-2i, 16-2i), the two bits of a pair swapped
    return ((r >> 1) & 0x5555u) | ((r & 0x5555u) << 1);
}

__global__ void __launch_bounds__(256) pl_pack(const unsigned char* __restrict__ bases, const u64* __restrict__ mm_start, const u8* __restrict__ mm_pos,
                                               const u64* __restrict__ pstart, const u8* __restrict__ size, int k, u64 n, u64 n_bases,
                                               u64* __restrict__ words)
{
    const u64 r = blockIdx.x * (u64)blockDim.x + threadIdx.x;
    if (r >= n) return;
    const unsigned char* const buf_end = bases + n_bases;
    const unsigned char* p = bases + (mm_start[r] - mm_pos[r]);       // first base of the super-k-mer
    int left = (int)size[r] + k - 1;                                     // bases to go
    u64 e = pstart[r];                                                   // element index in the packed stream
    u64 wi = e >> 5;
    int pos = (int)(e & 31);                                             // elements already in word wi
    u64 acc = 0;
    bool first_word = true;
    int skip = (int)(reinterpret_cast<uintptr_t>(p) & 7u);               // bytes of the first aligned word that lie before the record
    const u64* q = reinterpret_cast<const u64*>(p - skip);
    while (left > 0) {
        // an aligned word that sticks out of the buffer (its first or last few bytes) is put together byte by byte
        const unsigned char* g = reinterpret_cast<const unsigned char*>(q);
        u64 w8;
        if (g >= bases && g + 8 <= buf_end) w8 = __ldg(q);
        else {
            w8 = 0;
#pragma unroll 1
            for (int i = 0; i < 8; ++i) if (g + i >= bases && g + i < buf_end) w8 |= (u64)g[i] << (8 * i);
        }
        ++q;
        u32 chunk = codes8_msb_first(w8);
        int nb = 8 - skip;
        chunk = (chunk << (2 * skip)) & 0xFFFFu;
        skip = 0;
        if (nb > left) nb = left;
        chunk &= 0xFFFFu << (16 - 2 * nb);                               // codes of bytes behind the record out
        const u64 val = (u64)chunk << 48;                                // first base in the top two bits
        acc |= val >> (2 * pos);
        if (pos + nb >= 32) {
            if (first_word) atomicOr(reinterpret_cast<unsigned long long*>(words + wi), acc);   // shared with the previous record
            else words[wi] = acc;
            first_word = false;
            ++wi;
            acc = pos == 32 - nb ? 0 : val << (2 * (32 - pos));           // what did not fit
            pos = pos + nb - 32;
        } else pos += nb;
        left -= nb;
    }
    if (pos != 0) atomicOr(reinterpret_cast<unsigned long long*>(words + wi), acc);              // shared with the next record
}

cudaError_t scan_u64(u64* a, u64 n, u64* sums, u64* total, cudaStream_t stream)
{
    if (n == 0) return cudaMemsetAsync(total, 0, 8, stream);
    const u64 nb = (n + SC_BLOCK - 1) / SC_BLOCK;
    sc_sums<<<(unsigned)nb, SC_THREADS, 0, stream>>>(a, n, sums);
    sc_scan_sums<<<1, 1024, 0, stream>>>(sums, nb, total);
    sc_apply<<<(unsigned)nb, SC_THREADS, 0, stream>>>(a, n, sums);
    return cudaGetLastError();
}

}  // namespace

size_t payload_scratch_bytes(unsigned long long n_records, unsigned int n_buckets)
{
    const u64 n_tiles = (n_records + PL_TILE - 1) / PL_TILE + 1;
    const u64 m = n_tiles * n_buckets;
    const u64 big = m > n_records ? m : n_records;
    return (size_t)((m + (big + SC_BLOCK - 1) / SC_BLOCK + 16) * 8 + 256);
}

// All pointers device. words must be zeroed by the caller ((total bases * 2) / 64 + 1 entries, total bases <= n (w + k - 1)).
// o_pstart doubles as the length array before its scan. totals[0] = payload bases.
cudaError_t payload_launch(const unsigned char* bases, unsigned long long n_bases, const unsigned long long* value, const unsigned long long* mm_start, const unsigned char* size,
                           const unsigned char* mm_pos, unsigned long long n, int k, unsigned long long s8, unsigned int n_buckets, void* scratch,
                           unsigned long long* o_value, unsigned long long* o_mm, unsigned char* o_size, unsigned char* o_pos,
                           unsigned long long* o_pstart, unsigned long long* bucket_offsets, unsigned long long* words,
                           unsigned long long* totals, cudaStream_t stream, int* launches)
{
    *launches = 0;
    if (n == 0) {
        cudaError_t e = cudaMemsetAsync(bucket_offsets, 0, ((size_t)n_buckets + 1) * 8, stream);
        if (e != cudaSuccess) return e;
        return cudaMemsetAsync(totals, 0, 8, stream);
    }
    const u64 n_tiles = (n + PL_TILE - 1) / PL_TILE;
    const u64 m = n_tiles * n_buckets;
    u64* M = static_cast<u64*>(scratch);
    u64* sums = M + m;
    const size_t smem = (size_t)n_buckets * 4;
    pl_count<<<(unsigned)n_tiles, PL_THREADS, smem, stream>>>(value, n, s8, n_buckets, n_tiles, M);
    cudaError_t e = scan_u64(M, m, sums, totals + 1, stream);
    if (e != cudaSuccess) return e;
    pl_bucket_offsets<<<(n_buckets + 1 + 255) / 256, 256, 0, stream>>>(M, n_tiles, n_buckets, n, bucket_offsets);
    int bucket_bits = 0;
    while ((1u << bucket_bits) < n_buckets) ++bucket_bits;
    pl_scatter<<<(unsigned)n_tiles, PL_THREADS, (size_t)n_buckets * 12, stream>>>(value, mm_start, size, mm_pos, n, s8, n_buckets, bucket_bits, n_tiles, k, M, o_value, o_mm, o_size,
                                                                  o_pos, o_pstart);
    e = scan_u64(o_pstart, n, sums, totals, stream);
    if (e != cudaSuccess) return e;
    pl_pack<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(bases, o_mm, o_pos, o_pstart, o_size, k, n, n_bases, words);
    *launches = 10;
    return cudaGetLastError();
}

}  // namespace blg
