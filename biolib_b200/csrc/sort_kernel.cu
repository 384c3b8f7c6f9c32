// sort_kernel.cu -- the downstream pipeline of the reference's tests/test_jaccard.cpp:55-130 on the device:
//
//   emem::external_memory_vector<kmer_t>  (sorted runs + k-way merge = the k-mer values in sorted order,
//                                          reference include/external_memory_vector.hpp:186-196,244-251,320-341)
//   sampler::ordered_unique_sampler       (adjacent duplicates collapsed, include/ordered_unique_sampler.hpp:116-130)
//   algorithm::jaccard                    (merge of two sorted distinct streams: intersection / union / sizes, include/jaccard.hpp:8-38)
//
// as an LSD radix sort of the 64-bit k-mer values (8 bits per pass, only the ceil(2k / 8) passes the values can differ in), an
// adjacent-unique compaction, and |A n B| by looking the shorter distinct list up in the longer one (binary searches narrowed per block).
//
// One radix pass = per-tile digit histograms -> bucket-major exclusive scan -> stable scatter (rs_scatter: persistent CTAs, tiles
// staged by bulk async copies, warp-private ranking, tile reordered in shared memory, coalesced writes).
#include "common.cuh"
#include "kernels.h"

namespace blg {

namespace {

constexpr int RS_THREADS = 256, RS_WARPS = RS_THREADS / 32, RS_CHUNKS = 16, RS_TILE = RS_THREADS * RS_CHUNKS, RS_BINS = 256;

__global__ void __launch_bounds__(RS_THREADS) rs_count(const u64* __restrict__ keys, u64 n, int shift, u64 n_tiles, u64* __restrict__ M)
{
    __shared__ u32 hist[RS_BINS];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const u64 base = (u64)blockIdx.x * RS_TILE;
#pragma unroll 4
    for (int c = 0; c < RS_CHUNKS; ++c) {
        const u64 i = base + (u64)c * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&hist[(u32)(__ldg(keys + i) >> shift) & 255u], 1u);
    }
    __syncthreads();
    M[(u64)threadIdx.x * n_tiles + blockIdx.x] = hist[threadIdx.x];
}

// Stable scatter, persistent: a CTA walks tiles blockIdx.x, blockIdx.x + gridDim.x, ... The next tile's 32 KB of keys arrive by ONE
// bulk async copy (cp.async.bulk global -> shared, completion on an mbarrier; UBLKCP in SASS) while the current tile is ranked,
// reordered and written, so no warp ever waits for a key load (the first version loaded 16 keys per thread up front and then sat
// in long-scoreboard stalls for half of its time: DRAM 24 % busy). Ranking is warp-private (warp w owns keys [512 w, 512 w + 512)
// of the tile and walks them 32 at a time: rank = keys of the same digit seen by this warp so far + rank among the lanes of this
// round, from eight ballots), so the only block barriers are those between the phases; the tile is put in digit order in shared
// memory and written out with consecutive threads on consecutive addresses of each digit's run.
struct RsSmem {
    alignas(16) u64 in[2][RS_TILE];                // the tile as it arrives (double buffered)
    u64 skeys[RS_TILE];                            // the tile in digit order
    u64 gbase[RS_BINS];                            // global slot of a digit group's first key minus dstart
    u32 whist[RS_WARPS][RS_BINS];                  // keys per (warp, digit), then: first slot of (warp, digit) inside the tile's digit group
    u32 dstart[RS_BINS];                           // first slot of a digit's group inside the tile
    u32 wtot[RS_WARPS];
    alignas(8) unsigned long long mbar[2];
};

__device__ __forceinline__ u32 rs_smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void rs_mbar_init(unsigned long long* b, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(rs_smem_u32(b)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void rs_bulk_load(void* dst, const void* src, u32 bytes, unsigned long long* b)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(rs_smem_u32(b)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(rs_smem_u32(dst)), "l"(src), "r"(bytes), "r"(rs_smem_u32(b)) : "memory");
}
__device__ __forceinline__ void rs_mbar_wait(unsigned long long* b, u32 phase)
{
    asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}"
                 ::"r"(rs_smem_u32(b)), "r"(phase) : "memory");
}

__global__ void __launch_bounds__(RS_THREADS, 2) rs_scatter(const u64* __restrict__ keys, u64 n, int shift, u64 n_tiles, const u64* __restrict__ M,
                                                            u64* __restrict__ out)
{
    extern __shared__ __align__(16) unsigned char rs_raw[];
    RsSmem& sm = *reinterpret_cast<RsSmem*>(rs_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool bulk_ok = (reinterpret_cast<uintptr_t>(keys) & 15u) == 0;     // bulk copies want 16-byte aligned addresses and sizes
    if (tid == 0) { rs_mbar_init(&sm.mbar[0], 1); rs_mbar_init(&sm.mbar[1], 1); }
    __syncthreads();
    // a full tile of an aligned array comes by bulk copy; the last, partial tile (or an unaligned array) by ordinary loads
    auto stage = [&](u64 tile, int buf) -> bool {
        const u64 base = tile * RS_TILE;
        if (bulk_ok && base + RS_TILE <= n) {
            if (tid == 0) rs_bulk_load(sm.in[buf], keys + base, RS_TILE * 8, &sm.mbar[buf]);
            return true;
        }
        const u32 cnt = (u32)min((u64)RS_TILE, n - base);
        for (u32 i = tid; i < cnt; i += RS_THREADS) sm.in[buf][i] = __ldg(keys + base + i);
        return false;
    };
    u64 tile = blockIdx.x;
    if (tile >= n_tiles) return;
    bool cur_bulk = stage(tile, 0);
    u32 ph[2] = {0u, 0u};
    for (int it = 0; tile < n_tiles; ++it, tile += gridDim.x) {
        const int buf = it & 1;
        const u64 next = tile + gridDim.x;
        bool next_bulk = false;
        if (next < n_tiles) next_bulk = stage(next, buf ^ 1);       // in[buf ^ 1] was last read before the barrier that ended the previous iteration
        if (cur_bulk) { rs_mbar_wait(&sm.mbar[buf], ph[buf]); ph[buf] ^= 1u; }
        const u64 base = tile * RS_TILE;
        const u32 tile_n = (u32)min((u64)RS_TILE, n - base);
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) sm.whist[w][tid] = 0;
        __syncthreads();                                            // also: the ordinary loads of stage() are visible
        const u64* in = sm.in[buf];
        u32 rnk2[RS_CHUNKS / 2];                                    // ranks inside the warp's segment (< 512), two per register
        const u32 seg = (u32)warp * (RS_TILE / RS_WARPS);
#pragma unroll
        for (int r = 0; r < RS_CHUNKS; ++r) {
            const u32 li = seg + (u32)r * 32u + (u32)lane;
            const bool have = li < tile_n;
            const u32 d = have ? ((u32)(in[li] >> shift) & 255u) : 0u;
            u32 peers = __ballot_sync(0xffffffffu, have);
#pragma unroll
            for (int bit = 0; bit < 8; ++bit) {
                const u32 v = __ballot_sync(0xffffffffu, (d >> bit) & 1u);
                peers &= ((d >> bit) & 1u) ? v : ~v;
            }
            const u32 lr = __popc(peers & ((1u << lane) - 1u));
            const u32 prev = have ? sm.whist[warp][d] : 0u;
            if (r & 1) rnk2[r >> 1] |= (prev + lr) << 16; else rnk2[r >> 1] = prev + lr;
            __syncwarp();
            if (have && lr == 0) sm.whist[warp][d] = prev + __popc(peers);
            __syncwarp();
        }
        __syncthreads();
        {
            // digit tid: counts of the warps -> their first slots; tile count -> exclusive scan over the digits
            u32 run = 0;
#pragma unroll
            for (int w = 0; w < RS_WARPS; ++w) { const u32 x = sm.whist[w][tid]; sm.whist[w][tid] = run; run += x; }
            u32 incl = run;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const u32 v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
            if (lane == 31) sm.wtot[warp] = incl;
            __syncthreads();
            u32 before = 0;
#pragma unroll
            for (int w = 0; w < RS_WARPS; ++w) before += w < warp ? sm.wtot[w] : 0u;
            const u32 ds = before + incl - run;
            sm.dstart[tid] = ds;
            sm.gbase[tid] = M[(u64)tid * n_tiles + tile] - ds;
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < RS_CHUNKS; ++r) {
            const u32 li = seg + (u32)r * 32u + (u32)lane;
            if (li < tile_n) {
                const u64 key = in[li];
                const u32 d = (u32)(key >> shift) & 255u;
                sm.skeys[sm.dstart[d] + sm.whist[warp][d] + ((rnk2[r >> 1] >> (16 * (r & 1))) & 0xFFFFu)] = key;
            }
        }
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < RS_CHUNKS; ++j) {
            const u32 idx = (u32)j * RS_THREADS + (u32)tid;
            if (idx < tile_n) {
                const u64 k = sm.skeys[idx];
                out[sm.gbase[(u32)(k >> shift) & 255u] + idx] = k;
            }
        }
        __syncthreads();                                            // skeys / whist / in[buf] are free again
        cur_bulk = next_bulk;
    }
}

// the scatter's launch: persistent grid of 2 CTAs per SM (107 KB of shared memory each)
cudaError_t rs_scatter_launch(const u64* keys, u64 n, int shift, u64 n_tiles, const u64* M, u64* out, cudaStream_t stream)
{
    static int grid_cap[64] = {0};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (grid_cap[dev] == 0) {
        int sms = 0, occ = 0;
        if ((e = cudaFuncSetAttribute(rs_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RsSmem))) != cudaSuccess) return e;
        if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
        if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, rs_scatter, RS_THREADS, sizeof(RsSmem))) != cudaSuccess) return e;
        grid_cap[dev] = sms * (occ < 1 ? 1 : occ);
    }
    const unsigned grid = (unsigned)(n_tiles < (u64)grid_cap[dev] ? n_tiles : (u64)grid_cap[dev]);
    rs_scatter<<<grid, RS_THREADS, sizeof(RsSmem), stream>>>(keys, n, shift, n_tiles, M, out);
    return cudaGetLastError();
}

// ---- u64 exclusive scan (same three-kernel scheme as payload_kernel.cu) ---------------------------------------------------
constexpr int SC_THREADS = 256, SC_PER = 8, SC_BLOCK = SC_THREADS * SC_PER;

__global__ void __launch_bounds__(SC_THREADS) ss_sums(const u64* __restrict__ a, u64 n, u64* __restrict__ sums)
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
__global__ void __launch_bounds__(1024) ss_scan_sums(u64* __restrict__ sums, u64 n_blocks, u64* __restrict__ total)
{
    __shared__ u64 s[1024];
    const int tid = threadIdx.x;
    const u64 per = (n_blocks + 1023) / 1024;
    const u64 lo = min(n_blocks, (u64)tid * per), hi = min(n_blocks, lo + per);
    u64 t = 0;
    for (u64 i = lo; i < hi; ++i) t += sums[i];
    s[tid] = t;
    __syncthreads();
    if (tid == 0) { u64 r = 0; for (int i = 0; i < 1024; ++i) { const u64 x = s[i]; s[i] = r; r += x; } if (total) *total = r; }
    __syncthreads();
    u64 r = s[tid];
    for (u64 i = lo; i < hi; ++i) { const u64 x = sums[i]; sums[i] = r; r += x; }
}
__global__ void __launch_bounds__(SC_THREADS) ss_apply(u64* __restrict__ a, u64 n, const u64* __restrict__ sums)
{
    __shared__ u64 s[SC_THREADS];
    const u64 base = (u64)blockIdx.x * SC_BLOCK + (u64)threadIdx.x * SC_PER;
    u64 v[SC_PER], t = 0;
#pragma unroll
    for (int i = 0; i < SC_PER; ++i) { v[i] = base + i < n ? a[base + i] : 0; t += v[i]; }
    s[threadIdx.x] = t;
    __syncthreads();
    u64 r = sums[blockIdx.x];
    for (int j = 0; j < (int)threadIdx.x; ++j) r += s[j];
#pragma unroll
    for (int i = 0; i < SC_PER; ++i) { if (base + i < n) a[base + i] = r; r += v[i]; }
}
cudaError_t ss_scan(u64* a, u64 n, u64* sums, u64* total, cudaStream_t stream)
{
    if (n == 0) return total ? cudaMemsetAsync(total, 0, 8, stream) : cudaSuccess;
    const u64 nb = (n + SC_BLOCK - 1) / SC_BLOCK;
    ss_sums<<<(unsigned)nb, SC_THREADS, 0, stream>>>(a, n, sums);
    ss_scan_sums<<<1, 1024, 0, stream>>>(sums, nb, total);
    ss_apply<<<(unsigned)nb, SC_THREADS, 0, stream>>>(a, n, sums);
    return cudaGetLastError();
}

// ---- adjacent unique in two passes over the keys: per-block counts of run starts, then (after a scan of the few block counts)
// each block recomputes its flags and writes its distinct keys behind those of the blocks before it
constexpr int UQ_THREADS = 256, UQ_PER = 8, UQ_TILE = UQ_THREADS * UQ_PER;
template <bool WRITE>
__global__ void __launch_bounds__(UQ_THREADS) uq_tile(const u64* __restrict__ k, u64 n, u64* __restrict__ block_counts, u64* __restrict__ out)
{
    __shared__ u32 wsum[UQ_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u64 i0 = (u64)blockIdx.x * UQ_TILE + (u64)tid * UQ_PER;
    u64 v[UQ_PER];
    u64 prev = i0 > 0 && i0 <= n ? k[i0 - 1] : 0;
    u32 flags = 0;
#pragma unroll
    for (int j = 0; j < UQ_PER; ++j) {
        const u64 i = i0 + j;
        v[j] = i < n ? k[i] : 0;
        if (i < n && (i == 0 || v[j] != prev)) flags |= 1u << j;
        prev = v[j];
    }
    const u32 cnt = __popc(flags);
    u32 incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const u32 x = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += x; }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    u32 before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < UQ_THREADS / 32; ++w) { before += w < warp ? wsum[w] : 0u; total += wsum[w]; }
    if (!WRITE) {
        if (tid == 0) block_counts[blockIdx.x] = total;
        return;
    }
    u64 o = block_counts[blockIdx.x] + before + incl - cnt;     // block_counts: exclusive scan by now
#pragma unroll
    for (int j = 0; j < UQ_PER; ++j) if (flags & (1u << j)) out[o++] = v[j];
}

// ---- adjacent unique / generic keep-compaction ---------------------------------------------------------------------------------
// kmer_view's end test drops the k-mer that ends with the last byte of its sequence: keep[i] = 0 for those
__global__ void k1_flags(const u64* __restrict__ pos, u64 n, int k, const u64* __restrict__ offs, u64 n_seq, u64 n_bases, u64* __restrict__ f)
{
    const u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u64 end = pos[i] + (u64)k;
    u64 seq_end = n_bases;
    if (offs != nullptr) {
        u64 lo = 0, hi = n_seq;                    // first sequence boundary offs[q] (q >= 1) that is >= end
        while (lo < hi) { const u64 mid = (lo + hi) >> 1; if (offs[mid + 1] < end) lo = mid + 1; else hi = mid; }
        seq_end = lo < n_seq ? offs[lo + 1] : n_bases;
    }
    f[i] = end == seq_end ? 0 : 1;
}
// f_excl = exclusive scan of 0/1 flags, *total = their sum (device): element i is kept iff the scan steps at i
__global__ void compact_by_flags(const u64* __restrict__ k, const u64* __restrict__ f_excl, u64 n, const u64* __restrict__ total, u64* __restrict__ out)
{
    const u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u64 here = f_excl[i], next = i + 1 < n ? f_excl[i + 1] : *total;
    if (next != here) out[here] = k[i];
}

// ---- runs of consecutive super-k-mers (BLG_F_RUNS): record r opens a run iff it does not start where r-1 ended ----------------
__global__ void run_flags(const u64* __restrict__ mm_start, const u8* __restrict__ mm_pos, const u8* __restrict__ size, u64 n, u64* __restrict__ f)
{
    const u64 r = blockIdx.x * (u64)blockDim.x + threadIdx.x;
    if (r >= n) return;
    // mm_pos == nullptr: k-mer positions (blg_kmers with BLG_F_RUNS): every record starts at mm_start and is one position long
    if (mm_pos == nullptr) { f[r] = (r == 0 || mm_start[r] != mm_start[r - 1] + 1) ? 1 : 0; return; }
    const u64 sk = mm_start[r] - mm_pos[r];
    f[r] = (r == 0 || sk != mm_start[r - 1] - mm_pos[r - 1] + size[r - 1]) ? 1 : 0;
}
__global__ void run_scatter(const u64* __restrict__ mm_start, const u8* __restrict__ mm_pos, const u64* __restrict__ f_excl, u64 n,
                            const u64* __restrict__ total, u64* __restrict__ first, u64* __restrict__ start)
{
    const u64 r = blockIdx.x * (u64)blockDim.x + threadIdx.x;
    if (r >= n) return;
    const u64 here = f_excl[r], next = r + 1 < n ? f_excl[r + 1] : *total;
    if (next != here) { first[here] = r; start[here] = mm_start[r] - (mm_pos != nullptr ? mm_pos[r] : 0); }
}

// |A n B| of two sorted distinct arrays (algorithm::jaccard's merge, include/jaccard.hpp:8-38, as a count): a block takes 2048
// consecutive keys of A, narrows B to the range those keys can fall in (two binary searches), and every key is looked up in that
// range -- a dozen steps that stay in L1 -- instead of sorting the concatenation of the two lists a third time.
constexpr int IX_THREADS = 256, IX_PER = 8, IX_TILE = IX_THREADS * IX_PER;
__global__ void __launch_bounds__(IX_THREADS) intersect_count_kernel(const u64* __restrict__ a, u64 na, const u64* __restrict__ b, u64 nb,
                                                                     unsigned long long* __restrict__ count)
{
    __shared__ u64 range[2];
    __shared__ u32 wsum[IX_THREADS / 32];
    const u64 base = (u64)blockIdx.x * IX_TILE;
    const u64 last = min(base + IX_TILE, na) - 1;
    if (threadIdx.x < 2) {
        const u64 v = threadIdx.x == 0 ? a[base] : a[last];
        u64 lo = 0, hi = nb;                               // thread 0: first b >= a[base]; thread 1: first b > a[last]
        while (lo < hi) {
            const u64 mid = lo + (hi - lo) / 2;
            const u64 x = b[mid];
            if (threadIdx.x == 0 ? x < v : x <= v) lo = mid + 1; else hi = mid;
        }
        range[threadIdx.x] = lo;
    }
    __syncthreads();
    const u64 blo = range[0], bhi = range[1];
    u32 hits = 0;
#pragma unroll
    for (int c = 0; c < IX_PER; ++c) {
        const u64 i = base + (u64)c * IX_THREADS + threadIdx.x;
        if (i < na && bhi > blo) {
            const u64 v = a[i];
            u64 lo = blo, hi = bhi;
            while (lo < hi) {
                const u64 mid = lo + (hi - lo) / 2;
                if (b[mid] < v) lo = mid + 1; else hi = mid;
            }
            hits += (lo < bhi && b[lo] == v) ? 1u : 0u;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) hits += __shfl_xor_sync(0xffffffffu, hits, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = hits;
    __syncthreads();
    if (threadIdx.x == 0) {
        u32 t = 0;
        for (int w = 0; w < IX_THREADS / 32; ++w) t += wsum[w];
        if (t) atomicAdd(count, (unsigned long long)t);
    }
}

}  // namespace

// BLG_F_TIGHT: (value u64, size u8, mm_pos u8) -> (value bits 0..31, bits 32..39, (size - 1) | mm_pos << 4): 6 bytes per record over PCIe
namespace {
__global__ void __launch_bounds__(256) tight_pack_kernel(const u64* __restrict__ value, const u8* __restrict__ size, const u8* __restrict__ mm_pos, u64 n,
                                                         u32* __restrict__ lo32, u8* __restrict__ hi8, u8* __restrict__ size_pos)
{
    const u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u64 v = value[i];
    lo32[i] = (u32)v;
    if (hi8 != nullptr) hi8[i] = (u8)(v >> 32);
    size_pos[i] = (u8)(((u32)size[i] - 1u) | ((u32)mm_pos[i] << 4));
}
}  // namespace

cudaError_t tight_pack_launch(const unsigned long long* value, const unsigned char* size, const unsigned char* mm_pos, unsigned long long n,
                              unsigned int* lo32, unsigned char* hi8, unsigned char* size_pos, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    tight_pack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(value, size, mm_pos, n, lo32, hi8, size_pos);
    return cudaGetLastError();
}

size_t runs_scratch_bytes(unsigned long long n) { return (size_t)((3 * n + (n + SC_BLOCK - 1) / SC_BLOCK + 16) * 8 + 256); }

// scratch: [flags n][first n][start n][sums]; *d_total (device) = number of runs; first / start are returned as pointers into scratch
cudaError_t runs_launch(const unsigned long long* mm_start, const unsigned char* mm_pos, const unsigned char* size, unsigned long long n, void* scratch,
                        unsigned long long** first, unsigned long long** start, unsigned long long* d_total, cudaStream_t stream, int* launches)
{
    u64* flags = static_cast<u64*>(scratch);
    *first = flags + n;
    *start = flags + 2 * n;
    u64* sums = flags + 3 * n;
    if (n == 0) return cudaMemsetAsync(d_total, 0, 8, stream);
    const unsigned grid = (unsigned)((n + 255) / 256);
    run_flags<<<grid, 256, 0, stream>>>(mm_start, mm_pos, size, n, flags);
    cudaError_t e = ss_scan(flags, n, sums, d_total, stream);
    if (e != cudaSuccess) return e;
    run_scatter<<<grid, 256, 0, stream>>>(mm_start, mm_pos, flags, n, d_total, *first, *start);
    *launches += 5;
    return cudaGetLastError();
}

size_t sort_scratch_bytes(unsigned long long n)
{
    const u64 n_tiles = (n + RS_TILE - 1) / RS_TILE + 1;
    const u64 m = n_tiles * RS_BINS;
    const u64 big = m > n ? m : n;
    return (size_t)((m + (big + SC_BLOCK - 1) / SC_BLOCK + 16) * 8 + 256);
}

// In-place sort of keys[0, n) using tmp[0, n) as the second buffer; bits = number of low bits the keys can differ in.
cudaError_t radix_sort_u64(unsigned long long* keys, unsigned long long* tmp, unsigned long long n, int bits, void* scratch, cudaStream_t stream,
                           int* launches)
{
    if (n <= 1) return cudaSuccess;
    const u64 n_tiles = (n + RS_TILE - 1) / RS_TILE;
    const u64 m = n_tiles * RS_BINS;
    u64* M = static_cast<u64*>(scratch);
    u64* sums = M + m;
    u64 *src = keys, *dst = tmp;
    const int passes = (bits + 7) / 8;
    for (int p = 0; p < passes; ++p) {
        rs_count<<<(unsigned)n_tiles, RS_THREADS, 0, stream>>>(src, n, 8 * p, n_tiles, M);
        cudaError_t e = ss_scan(M, m, sums, nullptr, stream);
        if (e != cudaSuccess) return e;
        if ((e = rs_scatter_launch(src, n, 8 * p, n_tiles, M, dst, stream)) != cudaSuccess) return e;
        u64* t = src; src = dst; dst = t;
        *launches += 5;
    }
    if (src != keys) { cudaError_t e = cudaMemcpyAsync(keys, src, n * 8, cudaMemcpyDeviceToDevice, stream); if (e != cudaSuccess) return e; }
    return cudaGetLastError();
}

// keys sorted -> distinct keys in out (may not alias keys); *d_total (device) = their number. scratch: sort_scratch_bytes(n).
cudaError_t unique_sorted_u64(const unsigned long long* keys, unsigned long long n, unsigned long long* /*flags*/, void* scratch, unsigned long long* out,
                              unsigned long long* d_total, cudaStream_t stream, int* launches)
{
    if (n == 0) return cudaMemsetAsync(d_total, 0, 8, stream);
    const u64 nb = (n + UQ_TILE - 1) / UQ_TILE;
    u64* counts = static_cast<u64*>(scratch);
    u64* sums = counts + nb;
    uq_tile<false><<<(unsigned)nb, UQ_THREADS, 0, stream>>>(keys, n, counts, nullptr);
    cudaError_t e = ss_scan(counts, nb, sums, d_total, stream);
    if (e != cudaSuccess) return e;
    uq_tile<true><<<(unsigned)nb, UQ_THREADS, 0, stream>>>(keys, n, counts, out);
    *launches += 5;
    return cudaGetLastError();
}

// keep the k-mers that kmer_view yields (drop the last k-mer of every sequence that ends in a valid base): compacts `values` in place
// through tmp; *d_total = kept
cudaError_t drop_last_kmers(unsigned long long* values, const unsigned long long* pos, unsigned long long n, int k, const unsigned long long* offs,
                            unsigned long long n_seq, unsigned long long n_bases, unsigned long long* flags, void* scratch, unsigned long long* tmp,
                            unsigned long long* d_total, cudaStream_t stream, int* launches)
{
    if (n == 0) return cudaMemsetAsync(d_total, 0, 8, stream);
    const unsigned grid = (unsigned)((n + 255) / 256);
    k1_flags<<<grid, 256, 0, stream>>>(pos, n, k, offs, n_seq, n_bases, flags);
    cudaError_t e = ss_scan(flags, n, static_cast<u64*>(scratch), d_total, stream);
    if (e != cudaSuccess) return e;
    compact_by_flags<<<grid, 256, 0, stream>>>(values, flags, n, d_total, tmp);
    e = cudaMemcpyAsync(values, tmp, n * 8, cudaMemcpyDeviceToDevice, stream);
    *launches += 5;
    return e != cudaSuccess ? e : cudaGetLastError();
}

// index of the first key >= q[i] in a sorted array, one thread per query (the split points of a key-range partition: a handful of queries)
__global__ void lower_bound_kernel(const u64* __restrict__ keys, u64 n, const u64* __restrict__ q, u64 nq, u64* __restrict__ out)
{
    const u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const u64 v = q[i];
    u64 lo = 0, hi = n;
    while (lo < hi) {
        const u64 mid = lo + (hi - lo) / 2;
        if (keys[mid] < v) lo = mid + 1; else hi = mid;
    }
    out[i] = lo;
}

cudaError_t lower_bound_u64(const unsigned long long* keys, unsigned long long n, const unsigned long long* d_queries, unsigned long long nq,
                            unsigned long long* d_out, cudaStream_t stream)
{
    if (nq == 0) return cudaSuccess;
    lower_bound_kernel<<<(unsigned)((nq + 127) / 128), 128, 0, stream>>>(keys, n, d_queries, nq, d_out);
    return cudaGetLastError();
}

// *d_count (device) = |A n B| for two sorted arrays of distinct keys
cudaError_t intersect_count_u64(const unsigned long long* a, unsigned long long na, const unsigned long long* b, unsigned long long nb,
                                unsigned long long* d_count, cudaStream_t stream)
{
    cudaError_t e = cudaMemsetAsync(d_count, 0, 8, stream);
    if (e != cudaSuccess || na == 0 || nb == 0) return e;
    if (na > nb) { const unsigned long long* t = a; a = b; b = t; const unsigned long long tn = na; na = nb; nb = tn; }   // look the shorter list up in the longer
    intersect_count_kernel<<<(unsigned)((na + IX_TILE - 1) / IX_TILE), IX_THREADS, 0, stream>>>(a, na, b, nb, d_count);
    return cudaGetLastError();
}

// One stable scatter by the digit (key >> shift) & 255 (the partition step of the multi-GPU sort: keys grouped by their top byte,
// order inside a group kept). bin_start (device, 257 entries) = first slot of every digit's group, bin_start[256] = n.
cudaError_t partition_by_digit_u64(const unsigned long long* keys, unsigned long long* out, unsigned long long n, int shift, void* scratch,
                                   unsigned long long* bin_start, cudaStream_t stream, int* launches)
{
    if (n == 0) return cudaMemsetAsync(bin_start, 0, 257 * 8, stream);
    const u64 n_tiles = (n + RS_TILE - 1) / RS_TILE;
    const u64 m = n_tiles * RS_BINS;
    u64* M = static_cast<u64*>(scratch);
    u64* sums = M + m;
    rs_count<<<(unsigned)n_tiles, RS_THREADS, 0, stream>>>(keys, n, shift, n_tiles, M);
    cudaError_t e = ss_scan(M, m, sums, bin_start + 256, stream);          // total = n lands in bin_start[256]
    if (e != cudaSuccess) return e;
    e = cudaMemcpy2DAsync(bin_start, 8, M, n_tiles * 8, 8, 256, cudaMemcpyDeviceToDevice, stream);   // M[d][0]: where digit d starts
    if (e != cudaSuccess) return e;
    if ((e = rs_scatter_launch(keys, n, shift, n_tiles, M, out, stream)) != cudaSuccess) return e;
    *launches += 5;
    return cudaGetLastError();
}

}  // namespace blg
