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
//              later what is a base). Two passes over the text, each a grid of 16 KB blocks, with a one-CTA scan between:
//              (1) per block, for every state it may be entered in: kept bytes, record starts, state at its end; (2) the scan;
//              (3) with entry states and prefix sums known: compaction in shared memory, aligned 16-byte stores.
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
// Two passes over the text (grids of 16 KB blocks, 32 bytes per thread) with ONE single-CTA scan between them:
//   fx_summarise  per block, as a FUNCTION of the state the block is entered in: newlines (FASTQ: the state is the line number
//                 mod 4) / state of the line open at its end (FASTA), kept bytes and record starts for every possible entry state
//   fx_scan       entry state, first kept byte and first record of every block (one CTA, 1024 summaries per step)
//   fx_scatter    with all three known: kept bytes compacted in shared memory, written as whole 16-byte lines; record offsets
// Per thread everything is bit arithmetic on 32-bit masks (newline / '\r' / '>' found four bytes at a time); the only loops run
// over the newlines inside a thread's 32 bytes (0 or 1 for real files).
constexpr int FX_THREADS = 512;
constexpr int FX_BYTES_PER_THREAD = 32;
constexpr int FX_BLOCK = FX_THREADS * FX_BYTES_PER_THREAD;   // 16 KB of text per CTA
constexpr int FX_WARPS = FX_THREADS / 32;

// state of a line: FASTA 0 = sequence line, 1 = header line; FASTQ = line number mod 4 (1 = sequence line, 0 starts a record)
struct FxSum {
    u32 nl;                    // FASTQ: newlines in the block; FASTA: unused
    u32 last_state;            // FASTA: state of the line open at the block's end, 3 = no line starts in the block
    unsigned short keep[4];    // kept bytes if the block is entered in state e (FASTA: e = 0, 1)
    unsigned short starts[4];  // record starts if the block is entered in state e (FASTA: [0], the entry state does not matter)
};

// a thread's 32 bytes
struct Fx32 { uint4 a, b; };

__device__ __forceinline__ Fx32 load32(const unsigned char* text, size_t n, size_t pos)
{
    Fx32 v;
    if (pos + 32 <= n) {                                       // text is 16-byte aligned (checked by the caller)
        v.a = __ldg(reinterpret_cast<const uint4*>(text + pos));
        v.b = __ldg(reinterpret_cast<const uint4*>(text + pos) + 1);
        return v;
    }
    unsigned char t[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) t[i] = pos + i < n ? text[pos + i] : (unsigned char)'\n';   // the end of the text closes the last line
    memcpy(&v.a, t, 16);
    memcpy(&v.b, t + 16, 16);
    return v;
}

__device__ __forceinline__ unsigned char byte_of(const Fx32& v, int i)
{
    const uint4& q = i < 16 ? v.a : v.b;
    const int j = i & 15;
    const u32 w = j < 4 ? q.x : j < 8 ? q.y : j < 12 ? q.z : q.w;
    return (unsigned char)(w >> (8 * (j & 3)));
}

// bit q: byte q of w equals the byte replicated in pat (exact zero-byte test, then the four flags gathered by one multiply)
__device__ __forceinline__ u32 eqmask4(u32 w, u32 pat)
{
    const u32 x = w ^ pat;
    const u32 z = ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & 0x80808080u;
    return (((z >> 7) * 0x00204081u) >> 21) & 0xFu;
}
__device__ __forceinline__ u32 eqmask16(const uint4& v, u32 pat)
{
    return eqmask4(v.x, pat) | (eqmask4(v.y, pat) << 4) | (eqmask4(v.z, pat) << 8) | (eqmask4(v.w, pat) << 12);
}
__device__ __forceinline__ u32 eqmask32(const Fx32& v, u32 pat) { return eqmask16(v.a, pat) | (eqmask16(v.b, pat) << 16); }

// what a thread knows about its 32 bytes without any context
struct FxBits {
    u32 valid;   // bytes below n
    u32 nl;      // '\n'
    u32 body;    // neither '\n' nor '\r'
    u32 ls;      // first byte of a line
    u32 hdr;     // first byte of a line and '>'
};

__device__ __forceinline__ u32 below(int i) { return i >= 32 ? 0xFFFFFFFFu : ((1u << i) - 1u); }   // bits 0 .. i-1, i <= 32

__device__ __forceinline__ FxBits fx_bits(const unsigned char* text, size_t n, size_t base, const Fx32& v)
{
    FxBits b;
    b.valid = base >= n ? 0u : below(n - base >= 32 ? 32 : (int)(n - base));
    b.nl = eqmask32(v, 0x0A0A0A0Au) & b.valid;
    b.body = ~(b.nl | eqmask32(v, 0x0D0D0D0Du)) & b.valid;
    // is the thread's first byte the first byte of a line? the previous byte is the neighbour lane's last one
    const u32 pw = __shfl_up_sync(0xffffffffu, v.b.w, 1);
    u32 prev_nl;
    if ((threadIdx.x & 31) == 0) prev_nl = base == 0 || (base <= n && text[base - 1] == '\n');
    else prev_nl = (pw >> 24) == 0x0Au;
    b.ls = ((b.nl << 1) | prev_nl) & b.valid;
    b.hdr = b.ls & eqmask32(v, 0x3E3E3E3Eu);
    return b;
}

// FASTA: bit i set = byte i lies on a header line, given the state the thread is entered in
__device__ __forceinline__ u32 fasta_header_bytes(const FxBits& b, u32 entry)
{
    u32 in_hdr = entry ? 0xFFFFFFFFu : 0u;
    for (u32 s = b.ls; s; s &= s - 1) {
        const int p = __ffs((int)s) - 1;
        const u32 m = ~below(p);
        in_hdr = ((b.hdr >> p) & 1u) ? (in_hdr | m) : (in_hdr & ~m);
    }
    return in_hdr;
}
__device__ __forceinline__ u32 fasta_exit_state(const FxBits& b)   // 3: no line starts here, the entry state carries on
{
    return b.ls ? ((b.hdr >> (31 - __clz((int)b.ls))) & 1u) : 3u;
}

// FASTQ: keep / record-start masks, given the state the thread is entered in
__device__ __forceinline__ void fastq_masks(const FxBits& b, u32 state, u32* keep_out, u32* start_out)
{
    u32 keep = 0, start = (b.ls & 1u) && state == 0u ? 1u : 0u;
    int lo = 0;
    for (u32 s = b.nl; s; s &= s - 1) {
        const int p = __ffs((int)s) - 1;
        if (state == 1u) keep |= below(p) & ~below(lo);
        state = (state + 1u) & 3u;
        lo = p + 1;
        if (state == 0u && lo < 32) start |= 1u << lo;     // lo == 32 is the next thread's first byte
    }
    if (state == 1u) keep |= ~below(lo);
    *keep_out = keep & b.body;
    *start_out = start & b.valid;
}

// FASTQ without the entry state: counts by LOCAL phase (newlines seen so far mod 4), one byte per phase
__device__ __forceinline__ void fastq_phase_counts(const FxBits& b, u32* keep4, u32* start4)
{
    u32 kc = 0, sc = b.ls & 1u;
    int lo = 0, j = 0;
    for (u32 s = b.nl; s; s &= s - 1) {
        const int p = __ffs((int)s) - 1;
        kc += __popc(b.body & below(p) & ~below(lo)) << (8 * (j & 3));
        ++j;
        lo = p + 1;
        if (lo < 32 && ((b.valid >> lo) & 1u)) sc += 1u << (8 * (j & 3));
    }
    kc += __popc(b.body & ~below(lo)) << (8 * (j & 3));
    *keep4 = kc;
    *start4 = sc;
}

// sum over the warp of four byte-wide counters each (<= 32 per lane): two 16-bit fields per word, one REDUX per word
__device__ __forceinline__ void warp_sum4(u32 packed, u32* even, u32* odd)
{
    *even = __reduce_add_sync(0xffffffffu, packed & 0x00FF00FFu);           // counters 0 and 2
    *odd = __reduce_add_sync(0xffffffffu, (packed >> 8) & 0x00FF00FFu);     // counters 1 and 3
}

// exclusive prefix over the block of a per-thread count (<= 32 per field, two 16-bit fields); *total = the block's sum.
// The per-warp totals are scanned by every warp for itself (lane w holds warp w's total): no serial walk over the warps.
__device__ __forceinline__ u32 block_exclusive2(u32 x, u32* s_warp /* [FX_WARPS] */, u32* total)
{
    static_assert(FX_WARPS <= 32, "one lane per warp total");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u32 incl = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 y = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += y;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    u32 t = lane < FX_WARPS ? s_warp[lane] : 0u;
#pragma unroll
    for (int d = 1; d < FX_WARPS; d <<= 1) {
        const u32 y = __shfl_up_sync(0xffffffffu, t, d);
        if (lane >= d) t += y;
    }
    *total = __shfl_sync(0xffffffffu, t, FX_WARPS - 1);
    const u32 before = warp > 0 ? __shfl_sync(0xffffffffu, t, (warp + 31) & 31) : 0u;     // inclusive total of warp - 1
    return before + incl - x;
}

// FASTA: the state a thread is entered in = exit state of the nearest earlier thread of the block that has a line start.
// Returns it, or 2 when no earlier thread of the block has one (then it is the block's entry state). s_wst: [FX_WARPS].
// "Nearest earlier one that is defined" is the highest set bit below one's own position, within the warp (ballots over the
// lanes) and then over the warps (ballots over lanes that each hold one warp's summary).
__device__ __forceinline__ u32 fasta_thread_entry(u32 exit_state, u32* s_wst, u32* block_exit)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const u32 def = __ballot_sync(0xffffffffu, exit_state != 3u), hd = __ballot_sync(0xffffffffu, exit_state == 1u);
    if (lane == 0) s_wst[warp] = def ? ((hd >> (31 - __clz((int)def))) & 1u) : 3u;
    __syncthreads();
    const u32 ws = lane < FX_WARPS ? s_wst[lane] : 3u;
    const u32 wdef = __ballot_sync(0xffffffffu, ws != 3u), whd = __ballot_sync(0xffffffffu, ws == 1u);
    *block_exit = wdef ? ((whd >> (31 - __clz((int)wdef))) & 1u) : 3u;
    const u32 lower = def & ((1u << lane) - 1u);
    if (lower) return (hd >> (31 - __clz((int)lower))) & 1u;
    const u32 wlower = wdef & ((1u << warp) - 1u);
    if (wlower) return (whd >> (31 - __clz((int)wlower))) & 1u;
    return 2u;
}

__global__ void __launch_bounds__(FX_THREADS) fx_summarise(const unsigned char* __restrict__ text, size_t n, int fastq, FxSum* __restrict__ sum)
{
    __shared__ u32 s_a[FX_WARPS], s_b[FX_WARPS], s_c[FX_WARPS], s_d[FX_WARPS], s_e[FX_WARPS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t base = (size_t)blockIdx.x * FX_BLOCK + (size_t)tid * FX_BYTES_PER_THREAD;
    Fx32 v;
    v.a = v.b = make_uint4(0x0a0a0a0au, 0x0a0a0a0au, 0x0a0a0a0au, 0x0a0a0a0au);
    if (base < n) v = load32(text, n, base);
    const FxBits b = fx_bits(text, n, base, v);
    FxSum out;
    out.nl = 0; out.last_state = 3u;
#pragma unroll
    for (int q = 0; q < 4; ++q) { out.keep[q] = 0; out.starts[q] = 0; }
    if (fastq) {
        u32 total;
        const u32 pre = block_exclusive2(__popc(b.nl), s_a, &total);      // newlines of the block before this thread
        u32 k4, s4;
        fastq_phase_counts(b, &k4, &s4);
        // a byte of local phase j has state (E + pre + j) mod 4: rotate the counters to the block's phases
        const int rot = 8 * (int)(pre & 3u);
        k4 = __funnelshift_l(k4, k4, rot);
        s4 = __funnelshift_l(s4, s4, rot);
        u32 ke, ko, se, so;
        warp_sum4(k4, &ke, &ko);
        warp_sum4(s4, &se, &so);
        if (lane == 0) { s_b[warp] = ke; s_c[warp] = ko; s_d[warp] = se; s_e[warp] = so; }
        __syncthreads();
        if (tid == 0) {
            u32 a0 = 0, a1 = 0, a2 = 0, a3 = 0;
            for (int w = 0; w < FX_WARPS; ++w) { a0 += s_b[w]; a1 += s_c[w]; a2 += s_d[w]; a3 += s_e[w]; }
            const u32 kq[4] = {a0 & 0xFFFFu, a1 & 0xFFFFu, a0 >> 16, a1 >> 16};   // kept bytes whose state is (E + q) mod 4 ... kept iff that is 1
            const u32 sq[4] = {a2 & 0xFFFFu, a3 & 0xFFFFu, a2 >> 16, a3 >> 16};   // line starts whose state is (E + q) mod 4 ... records iff that is 0
            out.nl = total;
#pragma unroll
            for (int e = 0; e < 4; ++e) { out.keep[e] = (unsigned short)kq[(1 - e) & 3]; out.starts[e] = (unsigned short)sq[(4 - e) & 3]; }
            sum[blockIdx.x] = out;
        }
    } else {
        u32 block_exit;
        const u32 e = fasta_thread_entry(fasta_exit_state(b), s_a, &block_exit);
        // bytes before the thread's first line start follow the entry state, the others do not depend on it
        const int f = b.ls ? __ffs((int)b.ls) - 1 : 32;
        const u32 var = __popc(b.body & below(f));
        const u32 fix = __popc(~fasta_header_bytes(b, 0u) & b.body & ~below(f));
        const u32 sure = fix + (e == 0u ? var : 0u), open = e == 2u ? var : 0u;
        const u32 x = __reduce_add_sync(0xffffffffu, sure | (open << 16));          // <= 1024 per field and warp, <= 16384 per block
        const u32 st = __reduce_add_sync(0xffffffffu, (u32)__popc(b.hdr));
        if (lane == 0) { s_b[warp] = x; s_c[warp] = st; }
        __syncthreads();
        if (tid == 0) {
            u32 xs = 0, ss = 0;
            for (int w = 0; w < FX_WARPS; ++w) { xs += s_b[w]; ss += s_c[w]; }
            out.last_state = block_exit;
            out.keep[0] = (unsigned short)((xs & 0xFFFFu) + (xs >> 16));
            out.keep[1] = (unsigned short)(xs & 0xFFFFu);
            out.starts[0] = (unsigned short)ss;
            sum[blockIdx.x] = out;
        }
    }
}

// scan over the CTA's 1024 threads of an associative op whose identity is 0 (two shuffle levels); returns the inclusive value,
// *excl = the exclusive one
template <class Op>
__device__ __forceinline__ u64 cta_scan(u64 x, u64* s_w /* [32] */, Op op, u64* excl)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u64 y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x = op(y, x);
    }
    u64 before = __shfl_up_sync(0xffffffffu, x, 1);
    if (lane == 0) before = 0;
    __syncthreads();                 // s_w may still be read from the previous scan
    if (lane == 31) s_w[warp] = x;
    __syncthreads();
    if (warp == 0) {
        u64 t = s_w[lane];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const u64 y = __shfl_up_sync(0xffffffffu, t, d);
            if (lane >= d) t = op(y, t);
        }
        s_w[lane] = t;
    }
    __syncthreads();
    const u64 pre = warp > 0 ? s_w[warp - 1] : 0;
    *excl = op(pre, before);
    return op(pre, x);
}

// one CTA of 1024 threads walks the block summaries 1024 at a time (one per thread, coalesced): entry state, first kept byte
// and first record of every block; totals at index n_blocks and in *totals. The carried state is associative: FASTQ a sum
// mod 4, FASTA "the last defined value" (0 = undefined, 1 = sequence line, 2 = header line).
__global__ void __launch_bounds__(1024) fx_scan(const FxSum* __restrict__ sum, size_t n_blocks, int fastq, u32* __restrict__ entry,
                                                u64* __restrict__ keep_off, u64* __restrict__ start_off, unsigned long long* __restrict__ totals)
{
    __shared__ u64 s_w[32];
    __shared__ u64 s_carry[3];
    const int tid = threadIdx.x;
    const auto add = [](u64 a, u64 b) { return a + b; };
    const auto last = [](u64 a, u64 b) { return b ? b : a; };
    u64 c_state = 0, c_keep = 0, c_start = 0;       // over all earlier chunks (the text's first byte starts a line: no block reads the initial FASTA state)
    for (size_t chunk = 0; chunk < n_blocks; chunk += 1024) {
        const size_t b = chunk + tid;
        const bool have = b < n_blocks;
        FxSum x;
        x.nl = 0; x.last_state = 3u;
        if (have) x = sum[b];
        const u64 st = !have ? 0 : fastq ? x.nl : (x.last_state != 3u ? x.last_state + 1u : 0u);
        u64 ex;
        const u64 in = fastq ? cta_scan(st, s_w, add, &ex) : cta_scan(st, s_w, last, &ex);
        u32 state;
        if (fastq) state = (u32)((c_state + ex) & 3u);
        else { const u64 e = ex ? ex : c_state; state = e == 2u ? 1u : 0u; }
        const u64 kc = have ? x.keep[state] : 0, sc = have ? x.starts[fastq ? state : 0u] : 0;
        u64 kex, sex;
        const u64 kin = cta_scan(kc, s_w, add, &kex);
        const u64 sin = cta_scan(sc, s_w, add, &sex);
        if (have) { entry[b] = state; keep_off[b] = c_keep + kex; start_off[b] = c_start + sex; }
        if (tid == 1023) {
            s_carry[0] = fastq ? c_state + in : (in ? in : c_state);
            s_carry[1] = c_keep + kin;
            s_carry[2] = c_start + sin;
        }
        __syncthreads();
        c_state = s_carry[0]; c_keep = s_carry[1]; c_start = s_carry[2];
        __syncthreads();
    }
    if (tid == 0) {
        keep_off[n_blocks] = c_keep; start_off[n_blocks] = c_start;
        totals[0] = c_keep; totals[1] = c_start;          // mapped pinned host memory: read by the host after a stream synchronise
    }
}

// scatter: kept bytes go through shared memory at the byte alignment they will have in `bases`, so that every 16-byte chunk of
// the staging buffer is one aligned 16-byte chunk of the output
__global__ void __launch_bounds__(FX_THREADS) fx_scatter(const unsigned char* __restrict__ text, size_t n, int fastq, const u32* __restrict__ entry,
                                                         const u64* __restrict__ keep_off, const u64* __restrict__ start_off,
                                                         unsigned char* __restrict__ bases, u64* __restrict__ offsets, u64* __restrict__ first_off)
{
    __shared__ __align__(16) unsigned char s_out[FX_BLOCK + 32];
    __shared__ u32 s_a[FX_WARPS], s_b[FX_WARPS];
    const int tid = threadIdx.x;
    const size_t base = (size_t)blockIdx.x * FX_BLOCK + (size_t)tid * FX_BYTES_PER_THREAD;
    Fx32 v;
    v.a = v.b = make_uint4(0x0a0a0a0au, 0x0a0a0a0au, 0x0a0a0a0au, 0x0a0a0a0au);
    if (base < n) v = load32(text, n, base);
    const FxBits b = fx_bits(text, n, base, v);
    const u32 E = entry[blockIdx.x];
    u32 keep, start;
    if (fastq) {
        u32 total;
        const u32 pre = block_exclusive2(__popc(b.nl), s_a, &total);
        fastq_masks(b, (E + pre) & 3u, &keep, &start);
    } else {
        u32 block_exit;
        u32 e = fasta_thread_entry(fasta_exit_state(b), s_a, &block_exit);
        if (e == 2u) e = E;
        keep = ~fasta_header_bytes(b, e) & b.body;
        start = b.hdr;
    }
    u32 total;
    const u32 ex = block_exclusive2((u32)__popc(keep) | ((u32)__popc(start) << 16), s_b, &total);
    const u64 k0 = keep_off[blockIdx.x];
    const u32 kx = ex & 0xFFFFu, sx = ex >> 16, ktot = total & 0xFFFFu;
    unsigned char* dst = bases + k0;
    const u32 mis = (u32)(reinterpret_cast<uintptr_t>(dst) & 15u);
    {
        u32 p = mis + kx;
        if (keep == 0xFFFFFFFFu && (p & 3u) == 0u) {        // the usual thread of a sequence line, when the alignment allows words
            u32* w = reinterpret_cast<u32*>(s_out + p);
            w[0] = v.a.x; w[1] = v.a.y; w[2] = v.a.z; w[3] = v.a.w;
            w[4] = v.b.x; w[5] = v.b.y; w[6] = v.b.z; w[7] = v.b.w;
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if ((keep >> i) & 1u) s_out[p++] = byte_of(v, i);
        }
    }
    if (start) {
        u64 sp = start_off[blockIdx.x] + sx;
        for (u32 s = start; s; s &= s - 1) {
            const int i = __ffs((int)s) - 1;
            const u64 off = k0 + kx + __popc(keep & below(i));
            if (sp == 0) *first_off = off;                 // the caller checks that nothing is kept before the first record
            offsets[sp++] = off;
        }
    }
    if (blockIdx.x == 0 && tid == 0) offsets[start_off[gridDim.x]] = keep_off[gridDim.x];   // the closing entry: offsets[n_seq] = n_bases
    __syncthreads();
    const u32 end = mis + ktot;
    unsigned char* dst0 = dst - mis;                                   // 16-byte aligned
    for (u32 c = tid; c * 16u < end; c += FX_THREADS) {
        const u32 lo = c * 16u, hi = lo + 16u;
        if (lo >= mis && hi <= end) *reinterpret_cast<uint4*>(dst0 + lo) = *reinterpret_cast<const uint4*>(s_out + lo);
        else {
            const u32 a0 = lo < mis ? mis : lo, a1 = hi > end ? end : hi;
            for (u32 q = a0; q < a1; ++q) dst0[q] = s_out[q];
        }
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
    return (nb * (sizeof(FxSum) + sizeof(u32) + 2 * sizeof(u64)) + 511) / 256 * 256;
}

// Two stages, so that the caller can size the outputs exactly: fastx_count_launch summarises and scans and leaves
// totals[0] = kept bytes, totals[1] = records (totals: device address of mapped pinned host memory); fastx_emit_launch (same
// scratch, untouched in between) scatters into bases (totals[0] bytes) and offsets (totals[1] + 1 entries, the closing one
// included) and reports the first record's offset through *first_off (mapped pinned memory as well).
cudaError_t fastx_count_launch(const unsigned char* text, size_t n_text, int fastq, void* scratch, unsigned long long* totals, cudaStream_t stream,
                               int* launches)
{
    const size_t nb = (n_text + FX_BLOCK - 1) / FX_BLOCK;
    *launches = 0;
    if (nb == 0) return cudaErrorInvalidValue;     // the caller returns early for empty text
    char* p = static_cast<char*>(scratch);
    u64* keep = reinterpret_cast<u64*>(p); p += (nb + 1) * sizeof(u64);
    u64* starts = reinterpret_cast<u64*>(p); p += (nb + 1) * sizeof(u64);
    FxSum* sum = reinterpret_cast<FxSum*>(p); p += nb * sizeof(FxSum);
    u32* entry = reinterpret_cast<u32*>(p);
    fx_summarise<<<(unsigned)nb, FX_THREADS, 0, stream>>>(text, n_text, fastq, sum);
    fx_scan<<<1, 1024, 0, stream>>>(sum, nb, fastq, entry, keep, starts, totals);
    *launches = 2;
    return cudaGetLastError();
}

cudaError_t fastx_emit_launch(const unsigned char* text, size_t n_text, int fastq, void* scratch, unsigned char* bases, unsigned long long* offsets,
                              unsigned long long* first_off, cudaStream_t stream, int* launches)
{
    const size_t nb = (n_text + FX_BLOCK - 1) / FX_BLOCK;
    if (nb == 0) return cudaSuccess;
    char* p = static_cast<char*>(scratch);
    u64* keep = reinterpret_cast<u64*>(p); p += (nb + 1) * sizeof(u64);
    u64* starts = reinterpret_cast<u64*>(p); p += (nb + 1) * sizeof(u64);
    p += nb * sizeof(FxSum);
    u32* entry = reinterpret_cast<u32*>(p);
    fx_scatter<<<(unsigned)nb, FX_THREADS, 0, stream>>>(text, n_text, fastq, entry, keep, starts, bases, offsets, first_off);
    *launches += 1;
    return cudaGetLastError();
}

}  // namespace blg
