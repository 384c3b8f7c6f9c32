// ingest_kernel.cu -- input formats either side of the hot path, on the device:
//
//   unpack     2-bit packed bases (16 per 32-bit word, base q of a word at bits [2q, 2q+2): the little-endian stream the
//              fused kernels build internally, and the order of constants::seq_nt4_table / constants::bases, reference
//              include/constants.hpp:12-23) + a list of break intervals -> the ASCII buffer the kernels read. H2D then
//              moves 0.25 B per base instead of 1.
//   fastx      FASTA / FASTQ text -> concatenated sequence bytes + seq_offsets, what the reference gets record by record
//              from kseq_read() (reference tests/kseq.h:185-240, used by tests/test_kmer_view.cpp:30-40):
//                FASTA  a line whose first byte is '>' starts a record; every other line is sequence (multi-line allowed)
//                FASTQ  four lines per record: '@' header, sequence, '+' line, qualities
//              '\n' and '\r' are dropped, every other byte of a sequence line is kept as it is (the nt4 table decides
//              later what is a base). Three passes over the text, each a grid of 4 KB blocks, with one-CTA scans between:
//              (1) per block: newline count / state of its last line; (2) with the entry state known: kept bytes and
//              record starts per block; (3) with the prefix sums known: scatter.
#include "common.cuh"
#include "kernels.h"

namespace blg {

namespace {

// ---------------------------------------------------------------------------------------------------- unpack
__global__ void __launch_bounds__(256) unpack_kernel(const u32* __restrict__ codes, size_t n_words, uint4* __restrict__ out)
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n_words) return;
    const u32 c = __ldg(codes + i);
    // "ACGT"[code]: A 0x41, C 0x43, G 0x47, T 0x54 = 0x41 + {0, 2, 6, 0x13}
    u32 w[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        u32 r = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const u32 code = (c >> (8 * q + 2 * b)) & 3u;
            r |= ((0x54474341u >> (8 * code)) & 0xFFu) << (8 * b);
        }
        w[q] = r;
    }
    out[i] = make_uint4(w[0], w[1], w[2], w[3]);
}

__global__ void __launch_bounds__(128) stamp_breaks_kernel(unsigned char* __restrict__ out, size_t clip_lo, size_t clip_hi, const u64* __restrict__ start,
                                                           const u32* __restrict__ len, size_t n_breaks)
{
    // every break interval, clipped to the bases [clip_lo, clip_hi) that exist in `out` so far
    for (size_t b = blockIdx.x; b < n_breaks; b += gridDim.x) {
        const u64 s = start[b], e = s + len[b];
        const u64 lo = s > clip_lo ? s : clip_lo, hi = e < clip_hi ? e : clip_hi;
        for (u64 pos = lo + threadIdx.x; pos < hi; pos += blockDim.x) out[pos] = 'N';
    }
}

// ---------------------------------------------------------------------------------------------------- fastx
constexpr int FX_THREADS = 256;
constexpr int FX_BYTES_PER_THREAD = 16;
constexpr int FX_BLOCK = FX_THREADS * FX_BYTES_PER_THREAD;   // 4 KB of text per CTA

// state of a line: FASTA 0 = sequence line, 1 = header line; FASTQ = line number mod 4 (1 = sequence line)
struct FxSummary {
    u32 nl;          // newlines in the block
    u32 last_state;  // FASTA: state of the line that is open at the block's end, valid if nl > 0
};

__device__ __forceinline__ uint4 load16(const unsigned char* text, size_t n, size_t pos)
{
    if (pos + 16 <= n) return __ldg(reinterpret_cast<const uint4*>(text + pos));   // text is 16-byte aligned (checked by the caller)
    unsigned char b[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) b[i] = pos + i < n ? text[pos + i] : (unsigned char)'\n';   // the end of the text closes the last line
    uint4 v;
    memcpy(&v, b, 16);
    return v;
}

__device__ __forceinline__ unsigned char byte_of(const uint4& v, int i)
{
    const u32 w = i < 4 ? v.x : i < 8 ? v.y : i < 12 ? v.z : v.w;
    return (unsigned char)(w >> (8 * (i & 3)));
}

// pass 1: newline count of the block; FASTA: is the line open at the block's end a header line (known iff the block has a
// newline: the first byte after its last newline says)
__global__ void __launch_bounds__(FX_THREADS) fx_pass1(const unsigned char* __restrict__ text, size_t n, int fastq, FxSummary* __restrict__ sum)
{
    __shared__ u32 s_nl[FX_THREADS / 32];
    __shared__ int s_last[FX_THREADS / 32];
    const size_t base = (size_t)blockIdx.x * FX_BLOCK + (size_t)threadIdx.x * FX_BYTES_PER_THREAD;
    u32 nl = 0;
    int last = -1;        // offset (within the block) of the last newline seen by this thread
    if (base < n) {
        const uint4 v = load16(text, n, base);
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (base + i < n && byte_of(v, i) == '\n') { ++nl; last = threadIdx.x * FX_BYTES_PER_THREAD + i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { nl += __shfl_xor_sync(0xffffffffu, nl, o); last = max(last, __shfl_xor_sync(0xffffffffu, last, o)); }
    if ((threadIdx.x & 31) == 0) { s_nl[threadIdx.x >> 5] = nl; s_last[threadIdx.x >> 5] = last; }
    __syncthreads();
    if (threadIdx.x == 0) {
        u32 t = 0;
        int l = -1;
        for (int w = 0; w < FX_THREADS / 32; ++w) { t += s_nl[w]; l = max(l, s_last[w]); }
        FxSummary s;
        s.nl = t;
        s.last_state = 0;
        if (!fastq && l >= 0) {
            const size_t p = (size_t)blockIdx.x * FX_BLOCK + l + 1;     // first byte of the line open at the block's end (it may be
            s.last_state = p < n && text[p] == '>' ? 1u : 0u;           // the next block's first byte: same text)
        }
        sum[blockIdx.x] = s;
    }
}

// scan 1 (one CTA of 1024 threads): entry state of every block. The carried state is associative -- FASTQ: the line count mod 4
// (a sum); FASTA: the state after the last newline seen so far ("last defined value") -- so every thread summarises a contiguous
// run of blocks, the 1024 summaries are scanned in shared memory, and every thread walks its run again to write the entries.
__global__ void __launch_bounds__(1024) fx_scan1(const FxSummary* __restrict__ sum, size_t n_blocks, int fastq, const unsigned char* __restrict__ text,
                                                 size_t n, u32* __restrict__ entry)
{
    __shared__ u64 s_lines[1024];
    __shared__ u32 s_state[1024];                  // FASTA: state after the run's last newline, 3 = the run holds no newline
    const int tid = threadIdx.x;
    const size_t per = (n_blocks + 1023) / 1024;
    const size_t lo = min(n_blocks, (size_t)tid * per), hi = min(n_blocks, lo + per);
    u64 lines = 0;
    u32 st = 3u;
    for (size_t b = lo; b < hi; ++b) {
        const FxSummary x = sum[b];
        lines += x.nl;
        if (x.nl > 0) st = x.last_state;
    }
    s_lines[tid] = lines;
    s_state[tid] = st;
    __syncthreads();
    if (tid == 0) {
        // exclusive scan over the 1024 run summaries (shared memory only)
        u64 acc = 0;
        u32 state = fastq ? 0u : (n > 0 && text[0] == '>' ? 1u : 0u);     // the first line starts at byte 0
        for (int i = 0; i < 1024; ++i) {
            const u64 l = s_lines[i];
            const u32 t = s_state[i];
            s_lines[i] = acc;
            s_state[i] = state;
            acc += l;
            if (t != 3u) state = t;
        }
    }
    __syncthreads();
    u64 acc = s_lines[tid];
    u32 state = fastq ? (u32)(acc & 3u) : s_state[tid];
    for (size_t b = lo; b < hi; ++b) {
        entry[b] = state;
        const FxSummary x = sum[b];
        if (fastq) { acc += x.nl; state = (u32)(acc & 3u); }
        else if (x.nl > 0) state = x.last_state;
    }
}

// the per-thread walk shared by passes 2 and 3: state at the thread's first byte from a block-wide scan of (newline count,
// state of the last line) over the threads, then 16 bytes one by one. keep bit i: byte i is a sequence byte; start bit i: a
// record starts at byte i (first byte of a FASTA header line / of a FASTQ line with number mod 4 == 0).
struct FxWalk { u32 keep, start; };

__device__ __forceinline__ FxWalk fx_walk(const unsigned char* text, size_t n, int fastq, u32 block_entry, size_t block_first_byte_is_line_start,
                                          uint4* bytes_out)
{
    __shared__ u32 s_nl[FX_THREADS];
    __shared__ u32 s_state[FX_THREADS];     // FASTA: state after the thread's last newline (0/1), or 3 = no newline in the thread's bytes
    const int tid = threadIdx.x;
    const size_t base = (size_t)blockIdx.x * FX_BLOCK + (size_t)tid * FX_BYTES_PER_THREAD;
    uint4 v = make_uint4(0x0a0a0a0au, 0x0a0a0a0au, 0x0a0a0a0au, 0x0a0a0a0au);
    if (base < n) v = load16(text, n, base);
    *bytes_out = v;
    // thread summary
    u32 nl = 0, st = 3;
    {
        int last = -1;
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (base + i < n && byte_of(v, i) == '\n') { ++nl; last = i; }
        if (!fastq && last >= 0) {
            const size_t p = base + last + 1;
            st = p < n ? (text[p] == '>' ? 1u : 0u) : 0u;          // may read the next thread's / block's first byte: same text
        }
    }
    s_nl[tid] = nl;
    s_state[tid] = st;
    __syncthreads();
    // entry state of this thread: sequential look-back over the earlier threads of the block (256 at most; cheap next to the
    // global traffic, and only thread summaries are touched)
    u32 state = block_entry;
    if (fastq) {
        u32 lines = 0;
        for (int j = 0; j < tid; ++j) lines += s_nl[j];
        state = (block_entry + lines) & 3u;
    } else {
        for (int j = tid - 1; j >= 0; --j)
            if (s_state[j] != 3u) { state = s_state[j]; break; }
    }
    (void)block_first_byte_is_line_start;
    FxWalk w{0u, 0u};
    bool line_start = false;
    if (base < n) {
        // is the thread's first byte the first byte of a line?
        line_start = base == 0 || text[base - 1] == '\n';
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            if (base + i >= n) break;
            const unsigned char c = byte_of(v, i);
            if (line_start) {
                if (fastq) { if (state == 0u) w.start |= 1u << i; }
                else { state = c == '>' ? 1u : 0u; if (state == 1u) w.start |= 1u << i; }
                line_start = false;
            }
            if (c == '\n') { line_start = true; if (fastq) state = (state + 1u) & 3u; }
            else if (c != '\r' && (fastq ? state == 1u : state == 0u)) w.keep |= 1u << i;
        }
    }
    return w;
}

// pass 2: kept bytes and record starts of the block
__global__ void __launch_bounds__(FX_THREADS) fx_pass2(const unsigned char* __restrict__ text, size_t n, int fastq, const u32* __restrict__ entry,
                                                       u64* __restrict__ keep_cnt, u64* __restrict__ start_cnt)
{
    __shared__ u32 s_k[FX_THREADS / 32], s_s[FX_THREADS / 32];
    uint4 v;
    const FxWalk w = fx_walk(text, n, fastq, entry[blockIdx.x], 0, &v);
    u32 k = __popc(w.keep), s = __popc(w.start);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { k += __shfl_xor_sync(0xffffffffu, k, o); s += __shfl_xor_sync(0xffffffffu, s, o); }
    if ((threadIdx.x & 31) == 0) { s_k[threadIdx.x >> 5] = k; s_s[threadIdx.x >> 5] = s; }
    __syncthreads();
    if (threadIdx.x == 0) {
        u32 a = 0, b = 0;
        for (int q = 0; q < FX_THREADS / 32; ++q) { a += s_k[q]; b += s_s[q]; }
        keep_cnt[blockIdx.x] = a;
        start_cnt[blockIdx.x] = b;
    }
}

// scan 2 (one CTA of 1024 threads): exclusive prefix sums of the two per-block counts, totals at index n_blocks
__global__ void __launch_bounds__(1024) fx_scan2(u64* __restrict__ a, u64* __restrict__ b, size_t n_blocks)
{
    __shared__ u64 sa[1024], sb[1024];
    const int tid = threadIdx.x;
    const size_t per = (n_blocks + 1023) / 1024;
    const size_t lo = min(n_blocks, tid * per), hi = min(n_blocks, lo + per);
    u64 ta = 0, tb = 0;
    for (size_t i = lo; i < hi; ++i) { ta += a[i]; tb += b[i]; }
    sa[tid] = ta; sb[tid] = tb;
    __syncthreads();
    if (tid == 0) {
        u64 ra = 0, rb = 0;
        for (int i = 0; i < 1024; ++i) { const u64 xa = sa[i], xb = sb[i]; sa[i] = ra; sb[i] = rb; ra += xa; rb += xb; }
        a[n_blocks] = ra; b[n_blocks] = rb;
    }
    __syncthreads();
    u64 ra = sa[tid], rb = sb[tid];
    for (size_t i = lo; i < hi; ++i) { const u64 xa = a[i], xb = b[i]; a[i] = ra; b[i] = rb; ra += xa; rb += xb; }
}

// pass 3: scatter the kept bytes and the record offsets
__global__ void __launch_bounds__(FX_THREADS) fx_pass3(const unsigned char* __restrict__ text, size_t n, int fastq, const u32* __restrict__ entry,
                                                       const u64* __restrict__ keep_off, const u64* __restrict__ start_off,
                                                       unsigned char* __restrict__ bases, u64* __restrict__ offsets)
{
    __shared__ u32 s_k[FX_THREADS], s_s[FX_THREADS];
    uint4 v;
    const FxWalk w = fx_walk(text, n, fastq, entry[blockIdx.x], 0, &v);
    const int tid = threadIdx.x;
    __syncthreads();                       // fx_walk's shared arrays are done with
    s_k[tid] = __popc(w.keep);
    s_s[tid] = __popc(w.start);
    __syncthreads();
    u32 ko = 0, so = 0;
    for (int j = 0; j < tid; ++j) { ko += s_k[j]; so += s_s[j]; }
    u64 kp = keep_off[blockIdx.x] + ko, sp = start_off[blockIdx.x] + so;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (w.start & (1u << i)) offsets[sp++] = kp;
        if (w.keep & (1u << i)) bases[kp++] = byte_of(v, i);
    }
}

}  // namespace

// words [word_lo, word_hi) of the packed stream (codes points at word 0, on the device) -> bases [16 word_lo, 16 word_hi) of out, then the
// break intervals that fall inside them. unpack_launch does the whole input in one go; the chunked host path goes piece by piece.
cudaError_t unpack_range_launch(const void* codes, size_t word_lo, size_t word_hi, size_t n_bases, const unsigned long long* break_start,
                                const unsigned int* break_len, size_t n_breaks, unsigned char* out, cudaStream_t stream)
{
    if (word_hi <= word_lo) return cudaSuccess;
    const size_t nw = word_hi - word_lo;
    unpack_kernel<<<(unsigned)((nw + 255) / 256), 256, 0, stream>>>(static_cast<const u32*>(codes) + word_lo, nw, reinterpret_cast<uint4*>(out) + word_lo);
    if (n_breaks) {
        const unsigned grid = (unsigned)(n_breaks < 4096 ? n_breaks : 4096);
        const size_t hi = word_hi * 16 < n_bases ? word_hi * 16 : n_bases;
        stamp_breaks_kernel<<<grid, 128, 0, stream>>>(out, word_lo * 16, hi, break_start, break_len, n_breaks);
    }
    return cudaGetLastError();
}

cudaError_t unpack_launch(const void* codes, size_t n_bases, const unsigned long long* break_start, const unsigned int* break_len,
                          size_t n_breaks, unsigned char* out, cudaStream_t stream)
{
    return unpack_range_launch(codes, 0, (n_bases + 15) / 16, n_bases, break_start, break_len, n_breaks, out, stream);
}

size_t fastx_scratch_bytes(size_t n_text)
{
    const size_t nb = (n_text + FX_BLOCK - 1) / FX_BLOCK + 1;
    return (nb * (sizeof(FxSummary) + sizeof(u32) + 2 * sizeof(u64)) + 511) / 256 * 256;
}

// Two stages, so that the caller can size the outputs exactly: fastx_count_launch runs passes 1-2 and the scans and leaves
// totals[0] = kept bytes, totals[1] = records on the device; fastx_emit_launch (same scratch, untouched in between) scatters
// into bases (totals[0] bytes) and offsets (totals[1] entries; the caller appends the closing entry).
cudaError_t fastx_count_launch(const unsigned char* text, size_t n_text, int fastq, void* scratch, unsigned long long* totals, cudaStream_t stream,
                               int* launches)
{
    const size_t nb = (n_text + FX_BLOCK - 1) / FX_BLOCK;
    *launches = 0;
    if (nb == 0) return cudaMemsetAsync(totals, 0, 16, stream);
    char* p = static_cast<char*>(scratch);
    u64* keep = reinterpret_cast<u64*>(p); p += (nb + 1) * sizeof(u64);
    u64* starts = reinterpret_cast<u64*>(p); p += (nb + 1) * sizeof(u64);
    FxSummary* sum = reinterpret_cast<FxSummary*>(p); p += nb * sizeof(FxSummary);
    u32* entry = reinterpret_cast<u32*>(p);
    fx_pass1<<<(unsigned)nb, FX_THREADS, 0, stream>>>(text, n_text, fastq, sum);
    fx_scan1<<<1, 1024, 0, stream>>>(sum, nb, fastq, text, n_text, entry);
    fx_pass2<<<(unsigned)nb, FX_THREADS, 0, stream>>>(text, n_text, fastq, entry, keep, starts);
    fx_scan2<<<1, 1024, 0, stream>>>(keep, starts, nb);
    *launches = 4;
    cudaError_t e = cudaMemcpyAsync(totals, keep + nb, 8, cudaMemcpyDeviceToDevice, stream);
    if (e != cudaSuccess) return e;
    e = cudaMemcpyAsync(totals + 1, starts + nb, 8, cudaMemcpyDeviceToDevice, stream);
    if (e != cudaSuccess) return e;
    return cudaGetLastError();
}

cudaError_t fastx_emit_launch(const unsigned char* text, size_t n_text, int fastq, void* scratch, unsigned char* bases, unsigned long long* offsets,
                              cudaStream_t stream, int* launches)
{
    const size_t nb = (n_text + FX_BLOCK - 1) / FX_BLOCK;
    if (nb == 0) return cudaSuccess;
    char* p = static_cast<char*>(scratch);
    u64* keep = reinterpret_cast<u64*>(p); p += (nb + 1) * sizeof(u64);
    u64* starts = reinterpret_cast<u64*>(p); p += (nb + 1) * sizeof(u64);
    p += nb * sizeof(FxSummary);
    u32* entry = reinterpret_cast<u32*>(p);
    fx_pass3<<<(unsigned)nb, FX_THREADS, 0, stream>>>(text, n_text, fastq, entry, keep, starts, bases, offsets);
    *launches += 1;
    return cudaGetLastError();
}

}  // namespace blg
