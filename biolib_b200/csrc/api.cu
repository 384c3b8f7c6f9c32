// api.cu -- the C ABI declared in include/biolib_gpu.h: context, validation, transfers, launches.
// No algorithm lives here and nothing here computes on the CPU: without a usable CUDA device every
// compute entry point fails with BLG_E_CUDA.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/biolib_gpu.h"
#include "kernels.h"

using namespace blg;

namespace {

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};

constexpr int PIPE_MAX = 16;

struct Owner {
    int device_mem;  // 1: device memory, 0: pinned host memory
    void* ptrs[8];
    size_t bytes[8];
    blg_ctx* ctx;    // blocks go back to this context's pool
};

// Result arrays are handed out per call and given back through blg_free_*: cudaMallocHost / cudaMalloc of hundreds of MB cost
// more than the kernels (measured: 150 ms of a 210 ms shim call), so freed blocks are kept per context and reused.
struct BlockPool {
    struct Block { void* p; size_t bytes; };
    std::vector<Block> free_blocks;
    size_t cached = 0, limit = 0;
    bool device = false;
    void* take(size_t bytes, size_t* got)
    {
        int best = -1;
        for (int i = 0; i < (int)free_blocks.size(); ++i)
            if (free_blocks[i].bytes >= bytes && free_blocks[i].bytes <= 2 * bytes + (1u << 20) && (best < 0 || free_blocks[i].bytes < free_blocks[best].bytes)) best = i;
        if (best < 0) return nullptr;
        void* p = free_blocks[best].p;
        *got = free_blocks[best].bytes;
        cached -= *got;
        free_blocks.erase(free_blocks.begin() + best);
        return p;
    }
    // keeps the block unless it is larger than the whole pool; the blocks that have been lying here longest make room for it
    bool give(void* p, size_t bytes)
    {
        if (bytes > limit) return false;
        while (cached + bytes > limit && !free_blocks.empty()) {
            if (device) cudaFree(free_blocks.front().p); else cudaFreeHost(free_blocks.front().p);
            cached -= free_blocks.front().bytes;
            free_blocks.erase(free_blocks.begin());
        }
        free_blocks.push_back({p, bytes});
        cached += bytes;
        return true;
    }
};

}  // namespace

struct blg_ctx {
    int device = 0;
    int sm_count = 0;
    char name[128] = {0};
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaStream_t s_in = nullptr, s_out = nullptr;   // upload / download streams of the pipelined host path
    cudaEvent_t ev_in[16] = {};
    cudaEvent_t ev_dl[2] = {};                      // k-mer pipeline: the download of the piece that used device buffer b is done
    int pipe_first_div = 4;      // first pipeline piece = a regular piece / this
    int pipe_chunk_m = 12;       // k-mer starts per pipeline chunk, in Mi
    bool async_call = false;                 // BLG_F_ASYNC: run_sk_once only enqueues
    uint64_t async_capacity = 0;
    const unsigned char* staged = nullptr;  // set while a packed-input call runs: the bases are already unpacked in `in`
    const blg_packed_seq* packed_lazy = nullptr;   // ... or, for the chunked host path, still on the host: sk_pipelined uploads + unpacks piece by piece
    size_t lazy_breaks = 0;
    DevBuf packed, brk, fx_text, fx_scratch; // packed codes + break list of a packed-input call; FASTA/FASTQ text and parser scratch
    DevBuf in, offs, status, small;  // small: [0] tile counter (u32 in a u64 slot), [1] total, [2] tiles redone exactly (sk4)
    DevBuf o_value, o_mm, o_size, o_pos, o_hash, o_id, rank, o_lo, o_hi, o_position;
    DevBuf o_tlo, o_thi, o_tnib;            // BLG_F_TIGHT: packed value words / high bytes / size-position bytes
    cudaEvent_t ev_pk[16] = {};             // ... the pack kernel of pipeline piece i is done
    unsigned long long* h_total = nullptr;  // pinned + mapped: [0] record count, [1] cumulative exact-path tiles (sk4 writes both itself)
    unsigned long long* d_total = nullptr;  // the device's address of h_total
    // sk4: scan scratch that is never zeroed between launches (epoch-tagged words, a ticket counter that only grows)
    DevBuf sk4_status, sk4_small;           // sk4_small: [0] ticket counter, [2] cumulative exact-path tiles
    long long sk4_cap_tiles = 0;            // roles inside sk4_status are fixed per allocation: tiles | group accumulators | group words
    unsigned int sk4_epoch = 0, sk4_tickets = 0;
    unsigned long long sk4_exact_seen = 0;
    BlockPool pool_dev, pool_pinned;         // result arrays given back through blg_free_* (limits set in blg_create)
    std::string err;
    blg_stats stats{};
};

namespace {

#define CK(ctx, call)                                                                                  \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) {                                                                       \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                           \
            return e_ == cudaErrorMemoryAllocation ? BLG_E_NOMEM : BLG_E_CUDA;                         \
        }                                                                                              \
    } while (0)

int ensure(blg_ctx* c, DevBuf& b, size_t bytes)
{
    if (bytes <= b.cap) return BLG_OK;
    if (b.p) { CK(c, cudaFree(b.p)); b.p = nullptr; b.cap = 0; }
    size_t want = bytes + bytes / 8 + 256;
    CK(c, cudaMalloc(&b.p, want));
    b.cap = want;
    return BLG_OK;
}

struct Timer {
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    double ms() const { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); }
};

int check_offsets_host(const uint64_t* offs, size_t n_seq, size_t n_bases)
{
    if (!offs) return BLG_OK;
    if (n_seq == 0) return n_bases == 0 ? BLG_OK : BLG_E_INVALID;
    if (offs[0] != 0 || offs[n_seq] != n_bases) return BLG_E_INVALID;
    for (size_t i = 0; i < n_seq; ++i)
        if (offs[i] > offs[i + 1]) return BLG_E_INVALID;
    return BLG_OK;
}

// upload (or adopt) bases + offsets; returns device pointers
int place_input(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, uint32_t flags,
                const unsigned char** d_bases, const unsigned long long** d_offs)
{
    if (flags & BLG_F_DEVICE_INPUT) {
        *d_bases = reinterpret_cast<const unsigned char*>(bases);
        *d_offs = reinterpret_cast<const unsigned long long*>(offs);
        return BLG_OK;
    }
    int rc = check_offsets_host(offs, n_seq, n);
    if (rc != BLG_OK) { c->err = "seq_offsets must start at 0, end at n_bases and be non-decreasing"; return rc; }
    if (!c->staged) {
        if ((rc = ensure(c, c->in, n + 64)) != BLG_OK) return rc;
        if (n) CK(c, cudaMemcpyAsync(c->in.p, bases, n, cudaMemcpyHostToDevice, c->stream));
        c->stats.bytes_h2d += n;
    }
    *d_bases = static_cast<const unsigned char*>(c->in.p);
    *d_offs = nullptr;
    if (offs) {
        const size_t ob = (n_seq + 1) * sizeof(uint64_t);
        if ((rc = ensure(c, c->offs, ob)) != BLG_OK) return rc;
        CK(c, cudaMemcpyAsync(c->offs.p, offs, ob, cudaMemcpyHostToDevice, c->stream));
        c->stats.bytes_h2d += ob;
        *d_offs = static_cast<const unsigned long long*>(c->offs.p);
    }
    return BLG_OK;
}

int prep_scratch(blg_ctx* c, long long num_tiles)
{
    int rc;
    if ((rc = ensure(c, c->small, 64)) != BLG_OK) return rc;
    if ((rc = ensure(c, c->status, (size_t)(num_tiles > 0 ? num_tiles : 1) * 8)) != BLG_OK) return rc;
    CK(c, cudaMemsetAsync(c->small.p, 0, 64, c->stream));
    if (num_tiles > 0) CK(c, cudaMemsetAsync(c->status.p, 0, (size_t)num_tiles * 8, c->stream));
    return BLG_OK;
}

int read_total(blg_ctx* c, uint64_t* total)
{
    CK(c, cudaMemcpyAsync(c->h_total, static_cast<unsigned long long*>(c->small.p) + 1, 16, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    *total = c->h_total[0];
    c->stats.exact_tiles += c->h_total[1];
    c->stats.bytes_d2h += 16;
    return BLG_OK;
}

void add_kernel_time(blg_ctx* c)
{
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) == cudaSuccess) c->stats.kernel_ms += ms;
}

// one launch of the fused kernel into device arrays
int run_sk_once(blg_ctx* c, SkArgs& a, uint64_t* total)
{
    SkLaunchPlan plan;
    // one kernel per (k, m): sk4 for the benchmark families (k - m = 12, 10), the general form of sk3 for every other
    // k - m >= 5, sk2 below that
    const int gen = a.sync_on || sk4_supported(a.k, a.m) ? 4 : sk3_supported(a.k, a.m) ? 3 : 2;
    if (a.sync_on) CK(c, sy4_plan(a, c->sm_count, &plan));
    else if (gen == 4) CK(c, sk4_plan(a, c->sm_count, &plan));
    else if (gen == 2) CK(c, sk2_plan(a, c->sm_count, &plan));
    else CK(c, sk3_plan(a, c->sm_count, &plan));
    if (a.num_tiles == 0) { *total = 0; return BLG_OK; }
    int rc;
    if (gen == 4) {
        // no memset, no copy: see the protocol note in sk4_device.cuh
        if (a.num_tiles > c->sk4_cap_tiles || c->sk4_epoch >= (1u << 22) - 1 || !c->sk4_small.p) {
            const long long cap = a.num_tiles + a.num_tiles / 4 + 1024, groups = (cap + 31) / 32;
            if ((rc = ensure(c, c->sk4_status, (size_t)(cap + 2 * groups) * 8)) != BLG_OK) return rc;
            if ((rc = ensure(c, c->sk4_small, 64)) != BLG_OK) return rc;
            CK(c, cudaMemsetAsync(c->sk4_status.p, 0, c->sk4_status.cap, c->stream));
            CK(c, cudaMemsetAsync(c->sk4_small.p, 0, 64, c->stream));
            c->sk4_cap_tiles = cap; c->sk4_epoch = 0; c->sk4_tickets = 0; c->sk4_exact_seen = 0;
        }
        const long long cap_groups = (c->sk4_cap_tiles + 31) / 32;
        a.tile_counter = static_cast<unsigned int*>(c->sk4_small.p);
        a.exact_tiles = static_cast<unsigned long long*>(c->sk4_small.p) + 2;
        a.tile_status = static_cast<unsigned long long*>(c->sk4_status.p);
        a.grp_acc = a.tile_status + c->sk4_cap_tiles;
        a.grp_status = a.grp_acc + cap_groups;
        a.total = c->d_total;
        a.epoch = ++c->sk4_epoch;
        a.ticket_base = c->sk4_tickets;
        c->sk4_tickets += sk4_ticket_draws(a, plan);
        CK(c, cudaEventRecord(c->ev0, c->stream));
        if (a.sync_on) CK(c, sy4_launch(a, plan, c->stream)); else CK(c, sk4_launch(a, plan, c->stream));
        CK(c, cudaEventRecord(c->ev1, c->stream));
        c->stats.kernel_launches += 1;
        if (c->async_call) { *total = 0; c->async_capacity = a.capacity; return BLG_OK; }   // blg_async_count() collects
        CK(c, cudaStreamSynchronize(c->stream));
        *total = c->h_total[0];
        c->stats.exact_tiles += c->h_total[1] - c->sk4_exact_seen;
        c->sk4_exact_seen = c->h_total[1];
        add_kernel_time(c);
        return BLG_OK;
    }
    rc = prep_scratch(c, gen >= 3 ? a.status_words : a.num_tiles);
    if (rc != BLG_OK) return rc;
    a.tile_counter = static_cast<unsigned int*>(c->small.p);
    a.total = static_cast<unsigned long long*>(c->small.p) + 1;
    a.tile_status = static_cast<unsigned long long*>(c->status.p);
    if (gen >= 3) {
        const long long groups = (a.num_tiles + 31) / 32;
        a.grp_acc = a.tile_status + a.num_tiles;
        a.grp_status = a.grp_acc + groups;
    }
    CK(c, cudaEventRecord(c->ev0, c->stream));
    if (gen == 2) CK(c, sk2_launch(a, plan, c->stream));
    else CK(c, sk3_launch(a, plan, c->stream));
    CK(c, cudaEventRecord(c->ev1, c->stream));
    c->stats.kernel_launches += 1;
    rc = read_total(c, total);
    if (rc == BLG_OK) add_kernel_time(c);
    return rc;
}

int run_kmer_once(blg_ctx* c, KmerArgs& a, uint64_t* total)
{
    if (a.sync_on == 1 && sy4_supported(a.k, a.sync_m)) {
        // closed syncmers with k - m = 12 or 10: the filter in the sk4 frame (sy4_kernel.cu); kmer2_kernel<.., SYNC> has the rest
        SkArgs s{};
        s.bases = a.bases; s.buf_lo = a.buf_lo; s.buf_hi = a.buf_hi;
        s.seq_offsets = a.seq_offsets; s.n_seq = a.n_seq;
        s.win_lo = a.win_lo; s.win_hi = a.win_hi;
        s.k = a.k; s.m = a.sync_m; s.w = a.k - a.sync_m + 1;
        s.canonical = a.canonical; s.s8 = 8ULL; s.vec_ok = a.vec_ok;         // minimizer_position_extractor hashes with seed 0
        s.value = a.value_lo; s.mm_start = a.position; s.capacity = a.capacity;
        s.sync_on = 1; s.sync_s = a.sync_s; s.sync_e = a.sync_e;
        return run_sk_once(c, s, total);
    }
    KmerLaunchPlan plan;
    CK(c, kmer2_plan(a, c->sm_count, &plan));
    if (a.num_tiles == 0) { *total = 0; return BLG_OK; }
    int rc = prep_scratch(c, a.num_tiles);
    if (rc != BLG_OK) return rc;
    a.tile_counter = static_cast<unsigned int*>(c->small.p);
    a.total = static_cast<unsigned long long*>(c->small.p) + 1;
    a.tile_status = static_cast<unsigned long long*>(c->status.p);
    CK(c, cudaEventRecord(c->ev0, c->stream));
    CK(c, kmer2_launch(a, plan, c->stream));
    CK(c, cudaEventRecord(c->ev1, c->stream));
    c->stats.kernel_launches += 1;
    rc = read_total(c, total);
    if (rc == BLG_OK) add_kernel_time(c);
    return rc;
}

// BLG_F_EMIT_ID: rank directory over the valid-m-mer bits + one query per record (rank_kernel.cu)
int run_ids(blg_ctx* c, const SkArgs& a, uint64_t total, unsigned long long* d_id)
{
    if (total == 0) return BLG_OK;
    const long long n = a.buf_hi - a.buf_lo;
    int rc = ensure(c, c->rank, rank_scratch_bytes(n));
    if (rc != BLG_OK) return rc;
    int launches = 0;
    CK(c, record_ids_launch(a.bases, n, a.seq_offsets, a.n_seq, a.m, a.mm_start, total, c->rank.p, d_id, c->stream, &launches));
    c->stats.kernel_launches += launches;
    return BLG_OK;
}

int validate_sk(blg_ctx* c, const blg_params* p, const blg_sk_out* out)
{
    if (!c || !p || !out) return BLG_E_INVALID;
    if (p->k == 0 || p->k > 64 || p->m == 0 || p->m > 32 || p->m > p->k) {
        c->err = "need 1 <= m <= min(k, 32) and k <= 64";
        return BLG_E_INVALID;
    }
    if (p->flags & BLG_F_TIGHT) {
        if (!(p->flags & BLG_F_RUNS) || p->m > 20 || p->k - p->m > 15 || (p->flags & (BLG_F_EMIT_HASH | BLG_F_EMIT_ID))) {
            c->err = "BLG_F_TIGHT: with BLG_F_RUNS, m <= 20, k - m <= 15, no hash / id output";
            return BLG_E_INVALID;
        }
        if (out->capacity > 0 && (!out->value_lo32 || !out->size_pos || (p->m > 16 && !out->value_hi8))) { c->err = "BLG_F_TIGHT: capacity > 0 needs value_lo32, size_pos (and value_hi8 for m > 16)"; return BLG_E_INVALID; }
        return BLG_OK;
    }
    if (out->capacity > 0 && (!out->value || (!out->mm_start && !(p->flags & BLG_F_RUNS)))) { c->err = "capacity > 0 needs value and mm_start arrays"; return BLG_E_INVALID; }
    return BLG_OK;
}

uint64_t estimate_records(uint64_t nwin, int w)
{
    if (w <= 2) return nwin;
    // random hashes give 2/(w+1) records per k-mer; leave 30% + slack, never above one per k-mer
    const double est = (double)nwin * 2.0 / (w + 1) * 1.3 + 65536.0;
    return est > (double)nwin ? nwin : (uint64_t)est;
}

template <typename T>
int alloc_owned(blg_ctx* c, bool device_mem, size_t count, T** out, Owner* ow, int slot)
{
    size_t bytes = ((count ? count : 1) * sizeof(T) + 255) / 256 * 256;
    BlockPool& pool = device_mem ? c->pool_dev : c->pool_pinned;
    void* p = pool.take(bytes, &bytes);
    if (!p) {
        cudaError_t e = device_mem ? cudaMalloc(&p, bytes) : cudaMallocHost(&p, bytes);
        if (e == cudaErrorMemoryAllocation) {                 // give the cached blocks back to the driver and try once more
            cudaGetLastError();
            for (auto& b : pool.free_blocks) { if (device_mem) cudaFree(b.p); else cudaFreeHost(b.p); }
            pool.free_blocks.clear(); pool.cached = 0;
            e = device_mem ? cudaMalloc(&p, bytes) : cudaMallocHost(&p, bytes);
        }
        CK(c, e);
    }
    ow->ptrs[slot] = p;
    ow->bytes[slot] = bytes;
    ow->ctx = c;
    *out = static_cast<T*>(p);
    return BLG_OK;
}

void free_owner(Owner* ow)
{
    if (!ow) return;
    for (int i = 0; i < 8; ++i) {
        void* p = ow->ptrs[i];
        if (!p) continue;
        BlockPool* pool = ow->ctx ? (ow->device_mem ? &ow->ctx->pool_dev : &ow->ctx->pool_pinned) : nullptr;
        if (pool && pool->give(p, ow->bytes[i])) continue;
        if (ow->device_mem) cudaFree(p); else cudaFreeHost(p);
    }
    delete ow;
}

// Host in, host out, large input: the same work as the plain path below, cut into chunks of k-mer starts so that the
// upload of chunk i+1, the kernel of chunk i and the download of chunk i-1's records overlap (PCIe is full duplex and
// the record download is the longest leg). Chunks own disjoint window ranges of ONE device copy of the bases
// (blg_super_kmers_range semantics), so the concatenated output is the single-launch output.
// Returns BLG_OK with *done_flag = 1 when it produced the result, 0 when the caller should take the plain path
// (small input, id output, or the record estimate was too small).
int run_ids(blg_ctx* c, const SkArgs& a, uint64_t total, unsigned long long* d_id);

// BLG_F_RUNS: the run list of `total` device-resident records -> host arrays of `out` (library- or caller-owned)
int emit_runs(blg_ctx* c, const unsigned long long* d_mm, const unsigned char* d_pos, const unsigned char* d_size, uint64_t total, blg_sk_out* out,
              Owner* ow)
{
    out->n_runs = 0;
    if (total == 0) return BLG_OK;
    int rc = ensure(c, c->rank, runs_scratch_bytes(total) + 64);
    if (rc != BLG_OK) return rc;
    unsigned long long* d_tot = reinterpret_cast<unsigned long long*>(static_cast<char*>(c->rank.p) + runs_scratch_bytes(total) / 8 * 8);
    unsigned long long *d_first = nullptr, *d_start = nullptr, nr = 0;
    int launches = 0;
    CK(c, runs_launch(d_mm, d_pos, d_size, total, c->rank.p, &d_first, &d_start, d_tot, c->stream, &launches));
    c->stats.kernel_launches += launches;
    CK(c, cudaMemcpyAsync(&nr, d_tot, 8, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    out->n_runs = nr;
    if (ow) {
        if ((rc = alloc_owned(c, false, nr, &out->run_first_record, ow, 6)) != BLG_OK || (rc = alloc_owned(c, false, nr, &out->run_start, ow, 7)) != BLG_OK) return rc;
    } else if (nr > out->run_capacity || !out->run_first_record || !out->run_start) {
        c->err = "run arrays too small (out->n_runs holds the required size)";
        return BLG_E_CAPACITY;
    }
    CK(c, cudaMemcpyAsync(out->run_first_record, d_first, nr * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaMemcpyAsync(out->run_start, d_start, nr * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    c->stats.bytes_d2h += nr * 16 + 8;
    return BLG_OK;
}

// BLG_F_RUNS for blg_kmers: the run list of `total` device-resident k-mer positions, appended to two host vectors with the
// record indices shifted by `record_base` (the chunked path calls this once per piece)
int append_kmer_runs(blg_ctx* c, const unsigned long long* d_pos, uint64_t total, uint64_t record_base, std::vector<uint64_t>& first,
                     std::vector<uint64_t>& start)
{
    if (total == 0) return BLG_OK;
    int rc = ensure(c, c->rank, runs_scratch_bytes(total) + 64);
    if (rc != BLG_OK) return rc;
    unsigned long long* d_tot = reinterpret_cast<unsigned long long*>(static_cast<char*>(c->rank.p) + runs_scratch_bytes(total) / 8 * 8);
    unsigned long long *d_first = nullptr, *d_start = nullptr, nr = 0;
    int launches = 0;
    CK(c, runs_launch(d_pos, nullptr, nullptr, total, c->rank.p, &d_first, &d_start, d_tot, c->stream, &launches));
    c->stats.kernel_launches += launches;
    CK(c, cudaMemcpyAsync(&nr, d_tot, 8, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    const size_t at = first.size();
    first.resize(at + nr); start.resize(at + nr);
    CK(c, cudaMemcpyAsync(first.data() + at, d_first, nr * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaMemcpyAsync(start.data() + at, d_start, nr * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    c->stats.bytes_d2h += nr * 16 + 8;
    size_t w = at;
    for (size_t j = at; j < at + nr; ++j) {
        const uint64_t f = first[j] + record_base, s = start[j];
        // a piece boundary inside a stretch of valid bases: the run goes on
        if (w > 0 && j == at && start[w - 1] + (f - first[w - 1]) == s) continue;
        first[w] = f; start[w] = s; ++w;
    }
    first.resize(w); start.resize(w);
    return BLG_OK;
}

// hands the collected run list to the caller: library-owned pinned arrays (slots 3 and 4 of `ow`) or the caller's
int deliver_kmer_runs(blg_ctx* c, const std::vector<uint64_t>& first, const std::vector<uint64_t>& start, blg_kmer_out* out, Owner* ow)
{
    const uint64_t nr = first.size();
    out->n_runs = nr;
    if (ow) {
        int rc;
        if ((rc = alloc_owned(c, false, nr, &out->run_first_record, ow, 3)) != BLG_OK || (rc = alloc_owned(c, false, nr, &out->run_start, ow, 4)) != BLG_OK) return rc;
    } else if (nr > out->run_capacity || !out->run_first_record || !out->run_start) {
        c->err = "run arrays too small (out->n_runs holds the required size)";
        return BLG_E_CAPACITY;
    }
    if (nr) { memcpy(out->run_first_record, first.data(), nr * 8); memcpy(out->run_start, start.data(), nr * 8); }
    return BLG_OK;
}

int sk_pipelined(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, uint64_t lo, uint64_t hi,
                 const blg_params* p, blg_sk_out* out, bool with_size, bool want_hash, int* done_flag)
{
    *done_flag = 0;
    const bool want_id = (p->flags & BLG_F_EMIT_ID) != 0;
    const bool runs = (p->flags & BLG_F_RUNS) != 0;       // positions stay on the device; the run list is downloaded at the end
    const bool pos32 = (p->flags & BLG_F_POS32) != 0;
    const bool tight = (p->flags & BLG_F_TIGHT) != 0;  // records leave as (value_lo32, value_hi8, size_pos), packed per piece on the device
    const size_t posb = pos32 ? 4 : 8;                 // bytes per position entry
    const uint64_t nwin = hi - lo;
    const int k = p->k, m = p->m, w = k - m + 1;
    const bool thi = tight && m > 16;
    if (nwin < (8u << 20)) return BLG_OK;
    int rc;
    if ((rc = check_offsets_host(offs, n_seq, n)) != BLG_OK) { c->err = "seq_offsets must start at 0, end at n_bases and be non-decreasing"; return rc; }
    const uint64_t cap = out->capacity > 0 ? out->capacity : estimate_records(nwin, w);
    const blg_packed_seq* lazy = c->packed_lazy;        // packed input that is still on the host: uploaded and unpacked piece by piece below
    const bool staged = c->staged != nullptr && !lazy;  // packed input that is on the device already, unpacked: nothing to upload
    if (!staged && !lazy && (rc = ensure(c, c->in, n + 64)) != BLG_OK) return rc;
    if ((rc = ensure(c, c->o_value, cap * 8)) != BLG_OK || (rc = ensure(c, c->o_mm, cap * posb)) != BLG_OK) return rc;
    if (with_size && ((rc = ensure(c, c->o_size, cap)) != BLG_OK || (rc = ensure(c, c->o_pos, cap)) != BLG_OK)) return rc;
    if (want_hash && (rc = ensure(c, c->o_hash, cap * 8)) != BLG_OK) return rc;
    if (want_id && (rc = ensure(c, c->o_id, cap * 8)) != BLG_OK) return rc;
    if (tight && ((rc = ensure(c, c->o_tlo, cap * 4)) != BLG_OK || (rc = ensure(c, c->o_tnib, cap)) != BLG_OK || (thi && (rc = ensure(c, c->o_thi, cap)) != BLG_OK))) return rc;
    const unsigned long long* d_offs = nullptr;
    if (offs) {
        const size_t ob = (n_seq + 1) * sizeof(uint64_t);
        if ((rc = ensure(c, c->offs, ob)) != BLG_OK) return rc;
        CK(c, cudaMemcpyAsync(c->offs.p, offs, ob, cudaMemcpyHostToDevice, c->stream));
        c->stats.bytes_h2d += ob;
        d_offs = static_cast<const unsigned long long*>(c->offs.p);
    }
    // library-allocated host arrays are sized by the estimate; an overflow falls back to the plain path. The guard gives them
    // back (after draining the two side streams) on every early return below.
    Owner* ow = nullptr;
    struct Guard {
        blg_ctx* c; Owner** ow; bool armed;
        ~Guard() { if (armed) { cudaStreamSynchronize(c->s_in); cudaStreamSynchronize(c->s_out); cudaStreamSynchronize(c->stream); free_owner(*ow); *ow = nullptr; } }
    } guard{c, &ow, true};
    blg_sk_out h = *out;
    if (out->capacity == 0) {
        ow = new (std::nothrow) Owner{0, {}, {}, nullptr};
        if (!ow) return BLG_E_NOMEM;
        h.size = h.mm_pos = nullptr; h.hash = nullptr; h.id = nullptr; h.value = nullptr;
        h.value_lo32 = nullptr; h.value_hi8 = h.size_pos = nullptr;
        if (tight) {
            if ((rc = alloc_owned(c, false, cap, &h.value_lo32, ow, 0)) != BLG_OK || (rc = alloc_owned(c, false, cap, &h.size_pos, ow, 3)) != BLG_OK) return rc;
            if (thi && (rc = alloc_owned(c, false, cap, &h.value_hi8, ow, 2)) != BLG_OK) return rc;
        } else if ((rc = alloc_owned(c, false, cap, &h.value, ow, 0)) != BLG_OK) return rc;
        h.mm_start = nullptr;
        if (!runs && (rc = alloc_owned(c, false, pos32 ? (cap + 1) / 2 : cap, &h.mm_start, ow, 1)) != BLG_OK) return rc;
        if (with_size && !tight && ((rc = alloc_owned(c, false, cap, &h.size, ow, 2)) != BLG_OK || (rc = alloc_owned(c, false, cap, &h.mm_pos, ow, 3)) != BLG_OK)) return rc;
        if (want_hash && (rc = alloc_owned(c, false, cap, &h.hash, ow, 4)) != BLG_OK) return rc;
        if (want_id && (rc = alloc_owned(c, false, cap, &h.id, ow, 5)) != BLG_OK) return rc;
    }
    int nch = (int)(nwin / ((uint64_t)c->pipe_chunk_m << 20));
    nch = nch < 2 ? 2 : (nch > PIPE_MAX ? PIPE_MAX : nch);
    const size_t halo = (size_t)(k + w - 2) + 64;
    unsigned char* d_in = static_cast<unsigned char*>(c->in.p);
    uint64_t cut[PIPE_MAX + 1];
    size_t up[PIPE_MAX + 1];       // bases uploaded once piece i is in
    // the first piece is a fraction of the others: the record download (the longest of the three streams) starts sooner
    const uint64_t first = nwin / nch / (uint64_t)c->pipe_first_div;
    for (int i = 0; i <= nch; ++i) cut[i] = i == nch ? hi : i == 0 ? lo : (lo + first + (nwin - first) * (uint64_t)(i - 1) / (nch - 1)) / 16 * 16;
    size_t sent = 0;
    for (int i = 0; i < nch; ++i) {
        size_t end = i == nch - 1 ? n : (size_t)cut[i + 1] + halo;
        end = (end > n ? n : end);
        if (i < nch - 1) end = end / 16 * 16;
        if (staged) { up[i] = n; continue; }
        if (end > sent && lazy) {
            // 0.25 B per base over PCIe: the piece's packed words, then unpack + breaks on the upload stream (sent is a multiple of 16)
            const size_t w0 = sent / 16, w1 = (end + 15) / 16;
            CK(c, cudaMemcpyAsync(static_cast<char*>(c->packed.p) + w0 * 4, reinterpret_cast<const char*>(lazy->codes) + w0 * 4, (w1 - w0) * 4,
                                  cudaMemcpyHostToDevice, c->s_in));
            const unsigned long long* d_bs = static_cast<const unsigned long long*>(c->brk.p);
            CK(c, unpack_range_launch(c->packed.p, w0, w1, n, d_bs, reinterpret_cast<const unsigned int*>(d_bs + c->lazy_breaks), c->lazy_breaks, d_in, c->s_in));
            c->stats.kernel_launches += c->lazy_breaks ? 2 : 1;
        } else if (end > sent) CK(c, cudaMemcpyAsync(d_in + sent, bases + sent, end - sent, cudaMemcpyHostToDevice, c->s_in));
        CK(c, cudaEventRecord(c->ev_in[i], c->s_in));
        sent = end > sent ? end : sent;
        up[i] = sent;
    }
    if (!staged && !lazy) c->stats.bytes_h2d += n;
    uint64_t done = 0;
    bool overflow = false;
    for (int i = 0; i < nch && !overflow; ++i) {
        if (cut[i + 1] <= cut[i]) continue;
        if (!staged) CK(c, cudaStreamWaitEvent(c->stream, c->ev_in[i], 0));
        SkArgs a{};
        a.bases = d_in; a.buf_lo = 0; a.buf_hi = (long long)up[i];
        a.seq_offsets = d_offs; a.n_seq = (long long)n_seq;
        a.win_lo = (long long)cut[i]; a.win_hi = (long long)cut[i + 1];
        a.k = k; a.m = m; a.w = w;
        a.canonical = p->canonical ? 1 : 0;
        a.s8 = (unsigned long long)(uint32_t)p->seed ^ 8ULL;
        a.vec_ok = (reinterpret_cast<uintptr_t>(d_in) & 15u) == 0;
        a.value = static_cast<unsigned long long*>(c->o_value.p) + done;
        if (pos32) { a.mm_start32 = static_cast<unsigned int*>(c->o_mm.p) + done; a.pos_base = (long long)lo; }
        else a.mm_start = static_cast<unsigned long long*>(c->o_mm.p) + done;
        a.size = with_size ? static_cast<unsigned char*>(c->o_size.p) + done : nullptr;
        a.mm_pos = with_size ? static_cast<unsigned char*>(c->o_pos.p) + done : nullptr;
        a.hash = want_hash ? static_cast<unsigned long long*>(c->o_hash.p) + done : nullptr;
        a.capacity = cap - done;
        uint64_t cnt = 0;
        if ((rc = run_sk_once(c, a, &cnt)) != BLG_OK) return rc;   // returns after the kernel: count is known
        if (cnt > cap - done) { overflow = true; done += cnt; break; }
        if (cnt && tight) {
            // 6 bytes per record instead of 10: pack on the compute stream, download behind it
            unsigned int* t_lo = static_cast<unsigned int*>(c->o_tlo.p) + done;
            unsigned char* t_hi = thi ? static_cast<unsigned char*>(c->o_thi.p) + done : nullptr;
            unsigned char* t_nb = static_cast<unsigned char*>(c->o_tnib.p) + done;
            CK(c, tight_pack_launch(a.value, a.size, a.mm_pos, cnt, t_lo, t_hi, t_nb, c->stream));
            c->stats.kernel_launches += 1;
            CK(c, cudaEventRecord(c->ev_pk[i], c->stream));
            CK(c, cudaStreamWaitEvent(c->s_out, c->ev_pk[i], 0));
            CK(c, cudaMemcpyAsync(h.value_lo32 + done, t_lo, cnt * 4, cudaMemcpyDeviceToHost, c->s_out));
            if (thi) CK(c, cudaMemcpyAsync(h.value_hi8 + done, t_hi, cnt, cudaMemcpyDeviceToHost, c->s_out));
            CK(c, cudaMemcpyAsync(h.size_pos + done, t_nb, cnt, cudaMemcpyDeviceToHost, c->s_out));
            c->stats.bytes_d2h += cnt * (thi ? 6 : 5);
        } else if (cnt) {
            CK(c, cudaMemcpyAsync(h.value + done, a.value, cnt * 8, cudaMemcpyDeviceToHost, c->s_out));
            if (!runs) CK(c, cudaMemcpyAsync(reinterpret_cast<char*>(h.mm_start) + done * posb, static_cast<char*>(c->o_mm.p) + done * posb, cnt * posb, cudaMemcpyDeviceToHost, c->s_out));
            c->stats.bytes_d2h += cnt * (8 + (runs ? 0 : posb));
            if (with_size) {
                CK(c, cudaMemcpyAsync(h.size + done, a.size, cnt, cudaMemcpyDeviceToHost, c->s_out));
                CK(c, cudaMemcpyAsync(h.mm_pos + done, a.mm_pos, cnt, cudaMemcpyDeviceToHost, c->s_out));
                c->stats.bytes_d2h += cnt * 2;
            }
            if (want_hash) {
                CK(c, cudaMemcpyAsync(h.hash + done, a.hash, cnt * 8, cudaMemcpyDeviceToHost, c->s_out));
                c->stats.bytes_d2h += cnt * 8;
            }
        }
        done += cnt;
    }
    if (want_id && !overflow && done) {
        // minimizer_context_t::id: one rank-directory pass over the whole input once every chunk is in, then one more download
        SkArgs a{};
        a.bases = d_in; a.buf_lo = 0; a.buf_hi = (long long)n;
        a.seq_offsets = d_offs; a.n_seq = (long long)n_seq; a.m = m;
        a.mm_start = static_cast<unsigned long long*>(c->o_mm.p);
        if (!staged) CK(c, cudaStreamSynchronize(c->s_in));
        if ((rc = run_ids(c, a, done, static_cast<unsigned long long*>(c->o_id.p))) != BLG_OK) return rc;
        CK(c, cudaMemcpyAsync(h.id, c->o_id.p, done * 8, cudaMemcpyDeviceToHost, c->stream));
        CK(c, cudaStreamSynchronize(c->stream));
        c->stats.bytes_d2h += done * 8;
    }
    if (runs && !overflow) {
        if ((rc = emit_runs(c, static_cast<unsigned long long*>(c->o_mm.p), static_cast<unsigned char*>(c->o_pos.p), static_cast<unsigned char*>(c->o_size.p),
                            done, &h, out->capacity == 0 ? ow : nullptr)) != BLG_OK) {
            if (rc == BLG_E_CAPACITY) { out->n_runs = h.n_runs; out->count = done; }
            return rc;
        }
    }
    CK(c, cudaStreamSynchronize(c->s_out));
    if (!staged) CK(c, cudaStreamSynchronize(c->s_in));
    if (overflow) return BLG_OK;   // plain path: exact count, retry / BLG_E_CAPACITY (the guard releases the arrays)
    guard.armed = false;
    if (runs) { out->n_runs = h.n_runs; out->run_first_record = h.run_first_record; out->run_start = h.run_start; }
    if (out->capacity == 0) { *out = h; out->owner = ow; out->capacity = 0; }
    out->count = done;
    c->stats.n_records = done;
    *done_flag = 1;
    return BLG_OK;
}

int sk_common(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, uint64_t win_begin,
              uint64_t win_end, const blg_params* p, blg_sk_out* out, bool with_size)
{
    int rc = validate_sk(c, p, out);
    if (rc != BLG_OK) return rc;
    Timer tm;
    const uint64_t staged_h2d = c->staged ? c->stats.bytes_h2d : 0, staged_launches = c->staged ? c->stats.kernel_launches : 0;
    c->stats = blg_stats{};
    c->stats.n_bases = n;
    c->stats.bytes_h2d = staged_h2d;
    c->stats.kernel_launches = staged_launches;
    c->err.clear();
    CK(c, cudaSetDevice(c->device));
    const int k = p->k, m = p->m, w = k - m + 1;
    const bool dev_out = (p->flags & BLG_F_DEVICE_OUTPUT) != 0;
    const bool want_hash = (p->flags & BLG_F_EMIT_HASH) != 0;
    const bool want_id = (p->flags & BLG_F_EMIT_ID) != 0;
    const bool pos32 = (p->flags & BLG_F_POS32) != 0;
    const size_t posb = pos32 ? 4 : 8;
    if (pos32 && want_id) { c->err = "BLG_F_POS32 cannot be combined with BLG_F_EMIT_ID"; return BLG_E_INVALID; }
    const bool runs = (p->flags & BLG_F_RUNS) != 0;
    const bool tight = (p->flags & BLG_F_TIGHT) != 0, thi = tight && m > 16;
    if (runs && (dev_out || pos32 || !with_size || (p->flags & BLG_F_NO_SIZE))) { c->err = "BLG_F_RUNS: host output of the super-k-mer calls only, without BLG_F_POS32"; return BLG_E_INVALID; }
    with_size = with_size && !(p->flags & BLG_F_NO_SIZE);
    if (out->capacity > 0 && with_size && !tight && (!out->size || !out->mm_pos)) { c->err = "size / mm_pos arrays missing"; return BLG_E_INVALID; }
    if (out->capacity > 0 && want_hash && !out->hash) { c->err = "BLG_F_EMIT_HASH needs a hash array"; return BLG_E_INVALID; }
    if (out->capacity > 0 && want_id && !out->id) { c->err = "BLG_F_EMIT_ID needs an id array"; return BLG_E_INVALID; }

    uint64_t hi = n >= (size_t)k ? n - k + 1 : 0;
    if (win_end < hi) hi = win_end;
    const uint64_t lo = win_begin < hi ? win_begin : hi;
    const uint64_t nwin = hi - lo;
    out->count = 0;
    if (pos32 && nwin + (uint64_t)w > 0xFFFFFFFFull) { c->err = "BLG_F_POS32: the window range does not fit 32-bit positions"; return BLG_E_INVALID; }
    if (nwin == 0) {
        if (out->capacity == 0) {
            out->value = out->mm_start = out->hash = out->id = nullptr; out->size = out->mm_pos = nullptr; out->owner = nullptr;
            if (tight) { out->value_lo32 = nullptr; out->value_hi8 = out->size_pos = nullptr; }
            if (runs) { out->n_runs = 0; out->run_first_record = out->run_start = nullptr; }
        }
        c->stats.total_ms = tm.ms();
        return BLG_OK;
    }

    if (!(p->flags & (BLG_F_DEVICE_INPUT | BLG_F_DEVICE_OUTPUT))) {
        int produced = 0;
        if ((rc = sk_pipelined(c, bases, n, offs, n_seq, lo, hi, p, out, with_size, want_hash, &produced)) != BLG_OK) return rc;
        if (produced) { c->stats.total_ms = tm.ms(); return BLG_OK; }
    }

    const unsigned char* d_bases = nullptr;
    const unsigned long long* d_offs = nullptr;
    if ((rc = place_input(c, bases, n, offs, n_seq, p->flags, &d_bases, &d_offs)) != BLG_OK) return rc;

    SkArgs a{};
    a.bases = d_bases;
    a.buf_lo = 0; a.buf_hi = (long long)n;
    a.seq_offsets = d_offs; a.n_seq = (long long)n_seq;
    a.win_lo = (long long)lo; a.win_hi = (long long)hi;
    a.k = k; a.m = m; a.w = w;
    a.canonical = p->canonical ? 1 : 0;
    a.s8 = (unsigned long long)(uint32_t)p->seed ^ 8ULL;
    a.vec_ok = (reinterpret_cast<uintptr_t>(d_bases) & 15u) == 0;
    a.pos_base = (long long)lo;
    auto set_pos = [&](void* ptr) {        // the position array is uint32 (relative to lo) or uint64
        if (pos32) { a.mm_start32 = static_cast<unsigned int*>(ptr); a.mm_start = nullptr; }
        else { a.mm_start = static_cast<unsigned long long*>(ptr); a.mm_start32 = nullptr; }
    };

    uint64_t total = 0;
    c->async_call = (p->flags & BLG_F_ASYNC) != 0 && dev_out && (p->flags & BLG_F_DEVICE_INPUT) && out->capacity > 0 && !want_id &&
                    sk4_supported(k, m);
    struct AsyncOff { blg_ctx* c; ~AsyncOff() { c->async_call = false; } } async_off{c};
    if (dev_out && out->capacity > 0) {
        // caller-owned device arrays: one launch, no retry
        a.value = reinterpret_cast<unsigned long long*>(out->value);
        set_pos(out->mm_start);
        a.size = with_size ? out->size : nullptr;
        a.mm_pos = with_size ? out->mm_pos : nullptr;
        a.hash = want_hash ? reinterpret_cast<unsigned long long*>(out->hash) : nullptr;
        a.capacity = out->capacity;
        if ((rc = run_sk_once(c, a, &total)) != BLG_OK) return rc;
        out->count = total;
        c->stats.n_records = total;
        if (total > out->capacity) { c->stats.total_ms = tm.ms(); c->err = "output capacity too small"; return BLG_E_CAPACITY; }
        if (want_id && (rc = run_ids(c, a, total, reinterpret_cast<unsigned long long*>(out->id))) != BLG_OK) return rc;
        c->stats.total_ms = tm.ms();
        return BLG_OK;
    }

    // library-managed device arrays (grow-only scratch for host output, fresh allocations for device output)
    uint64_t cap = estimate_records(nwin, w);
    if (!dev_out && out->capacity > 0 && out->capacity < cap) cap = out->capacity;
    Owner* ow = nullptr;
    unsigned long long* d_id = nullptr;
    for (int attempt = 0; attempt < 2; ++attempt) {
        if (dev_out) {
            free_owner(ow);
            ow = new (std::nothrow) Owner{1, {}, {}, nullptr};
            if (!ow) return BLG_E_NOMEM;
            unsigned long long *v = nullptr, *s = nullptr, *h = nullptr, *idp = nullptr;
            unsigned char *sz = nullptr, *mp = nullptr;
            if ((rc = alloc_owned(c, true, cap, &v, ow, 0)) != BLG_OK || (rc = alloc_owned(c, true, pos32 ? (cap + 1) / 2 : cap, &s, ow, 1)) != BLG_OK) { free_owner(ow); return rc; }
            if (with_size && ((rc = alloc_owned(c, true, cap, &sz, ow, 2)) != BLG_OK || (rc = alloc_owned(c, true, cap, &mp, ow, 3)) != BLG_OK)) { free_owner(ow); return rc; }
            if (want_hash && (rc = alloc_owned(c, true, cap, &h, ow, 4)) != BLG_OK) { free_owner(ow); return rc; }
            if (want_id && (rc = alloc_owned(c, true, cap, &idp, ow, 5)) != BLG_OK) { free_owner(ow); return rc; }
            a.value = v; set_pos(s); a.size = sz; a.mm_pos = mp; a.hash = h; d_id = idp;
        } else {
            if ((rc = ensure(c, c->o_value, cap * 8)) != BLG_OK || (rc = ensure(c, c->o_mm, cap * posb)) != BLG_OK) return rc;
            if (with_size && ((rc = ensure(c, c->o_size, cap)) != BLG_OK || (rc = ensure(c, c->o_pos, cap)) != BLG_OK)) return rc;
            if (want_hash && (rc = ensure(c, c->o_hash, cap * 8)) != BLG_OK) return rc;
            if (want_id && (rc = ensure(c, c->o_id, cap * 8)) != BLG_OK) return rc;
            d_id = want_id ? static_cast<unsigned long long*>(c->o_id.p) : nullptr;
            a.value = static_cast<unsigned long long*>(c->o_value.p);
            set_pos(c->o_mm.p);
            a.size = with_size ? static_cast<unsigned char*>(c->o_size.p) : nullptr;
            a.mm_pos = with_size ? static_cast<unsigned char*>(c->o_pos.p) : nullptr;
            a.hash = want_hash ? static_cast<unsigned long long*>(c->o_hash.p) : nullptr;
        }
        a.capacity = cap;
        if ((rc = run_sk_once(c, a, &total)) != BLG_OK) { free_owner(ow); return rc; }
        if (total <= cap) break;
        if (!dev_out && out->capacity > 0 && total > out->capacity) break;  // the caller's arrays cannot hold it anyway
        cap = total;                                                       // exact size, run again
    }
    out->count = total;
    c->stats.n_records = total;
    const bool fits = total <= cap;
    if (want_id && fits && (rc = run_ids(c, a, total, d_id)) != BLG_OK) { free_owner(ow); return rc; }
    if (dev_out) {
        out->id = reinterpret_cast<uint64_t*>(d_id);
        out->value = reinterpret_cast<uint64_t*>(a.value);
        out->mm_start = pos32 ? reinterpret_cast<uint64_t*>(a.mm_start32) : reinterpret_cast<uint64_t*>(a.mm_start);
        out->size = a.size; out->mm_pos = a.mm_pos;
        out->hash = reinterpret_cast<uint64_t*>(a.hash);
        out->owner = ow;
        c->stats.total_ms = tm.ms();
        return BLG_OK;
    }
    if (out->capacity > 0) {
        if (total > out->capacity) { c->stats.total_ms = tm.ms(); c->err = "output capacity too small"; return BLG_E_CAPACITY; }
    } else {
        ow = new (std::nothrow) Owner{0, {}, {}, nullptr};
        if (!ow) return BLG_E_NOMEM;
        out->size = out->mm_pos = nullptr; out->hash = nullptr; out->id = nullptr; out->value = nullptr;
        if (tight) {
            out->value_hi8 = nullptr;
            if ((rc = alloc_owned(c, false, total, &out->value_lo32, ow, 0)) != BLG_OK || (rc = alloc_owned(c, false, total, &out->size_pos, ow, 3)) != BLG_OK ||
                (thi && (rc = alloc_owned(c, false, total, &out->value_hi8, ow, 2)) != BLG_OK)) { free_owner(ow); return rc; }
        } else if ((rc = alloc_owned(c, false, total, &out->value, ow, 0)) != BLG_OK) { free_owner(ow); return rc; }
        out->mm_start = nullptr;
        if (!runs && (rc = alloc_owned(c, false, pos32 ? (total + 1) / 2 : total, &out->mm_start, ow, 1)) != BLG_OK) { free_owner(ow); return rc; }
        if (with_size && !tight && ((rc = alloc_owned(c, false, total, &out->size, ow, 2)) != BLG_OK || (rc = alloc_owned(c, false, total, &out->mm_pos, ow, 3)) != BLG_OK)) { free_owner(ow); return rc; }
        if (want_hash && (rc = alloc_owned(c, false, total, &out->hash, ow, 4)) != BLG_OK) { free_owner(ow); return rc; }
        if (want_id && (rc = alloc_owned(c, false, total, &out->id, ow, 5)) != BLG_OK) { free_owner(ow); return rc; }
        out->owner = ow;
    }
    if (total && tight) {
        if ((rc = ensure(c, c->o_tlo, total * 4)) != BLG_OK || (rc = ensure(c, c->o_tnib, total)) != BLG_OK || (thi && (rc = ensure(c, c->o_thi, total)) != BLG_OK)) {
            if (out->capacity == 0) { free_owner(ow); out->owner = nullptr; }
            return rc;
        }
        unsigned char* t_hi = thi ? static_cast<unsigned char*>(c->o_thi.p) : nullptr;
        CK(c, tight_pack_launch(a.value, a.size, a.mm_pos, total, static_cast<unsigned int*>(c->o_tlo.p), t_hi, static_cast<unsigned char*>(c->o_tnib.p), c->stream));
        c->stats.kernel_launches += 1;
        CK(c, cudaMemcpyAsync(out->value_lo32, c->o_tlo.p, total * 4, cudaMemcpyDeviceToHost, c->stream));
        if (thi) CK(c, cudaMemcpyAsync(out->value_hi8, t_hi, total, cudaMemcpyDeviceToHost, c->stream));
        CK(c, cudaMemcpyAsync(out->size_pos, c->o_tnib.p, total, cudaMemcpyDeviceToHost, c->stream));
        c->stats.bytes_d2h += total * (thi ? 6 : 5);
        CK(c, cudaStreamSynchronize(c->stream));
        if ((rc = emit_runs(c, a.mm_start, a.mm_pos, a.size, total, out, out->capacity == 0 ? ow : nullptr)) != BLG_OK) {
            if (out->capacity == 0) { free_owner(ow); out->owner = nullptr; }
            return rc;
        }
    } else if (total) {
        CK(c, cudaMemcpyAsync(out->value, a.value, total * 8, cudaMemcpyDeviceToHost, c->stream));
        if (!runs) CK(c, cudaMemcpyAsync(out->mm_start, pos32 ? static_cast<void*>(a.mm_start32) : static_cast<void*>(a.mm_start), total * posb, cudaMemcpyDeviceToHost, c->stream));
        c->stats.bytes_d2h += total * (8 + (runs ? 0 : posb));
        if (with_size) {
            CK(c, cudaMemcpyAsync(out->size, a.size, total, cudaMemcpyDeviceToHost, c->stream));
            CK(c, cudaMemcpyAsync(out->mm_pos, a.mm_pos, total, cudaMemcpyDeviceToHost, c->stream));
            c->stats.bytes_d2h += total * 2;
        }
        if (want_hash) {
            CK(c, cudaMemcpyAsync(out->hash, a.hash, total * 8, cudaMemcpyDeviceToHost, c->stream));
            c->stats.bytes_d2h += total * 8;
        }
        if (want_id) {
            CK(c, cudaMemcpyAsync(out->id, d_id, total * 8, cudaMemcpyDeviceToHost, c->stream));
            c->stats.bytes_d2h += total * 8;
        }
        CK(c, cudaStreamSynchronize(c->stream));
        if (runs && (rc = emit_runs(c, a.mm_start, a.mm_pos, a.size, total, out, out->capacity == 0 ? ow : nullptr)) != BLG_OK) {
            if (out->capacity == 0) { free_owner(ow); out->owner = nullptr; }
            return rc;
        }
    }
    c->stats.total_ms = tm.ms();
    return BLG_OK;
}

}  // namespace

extern "C" {

int blg_abi_version(void) { return BLG_ABI_VERSION; }

const char* blg_strerror(int status)
{
    switch (status) {
    case BLG_OK: return "ok";
    case BLG_E_INVALID: return "invalid argument";
    case BLG_E_UNSUPPORTED: return "unsupported combination";
    case BLG_E_CAPACITY: return "output capacity too small";
    case BLG_E_CUDA: return "CUDA error (no device, or a runtime failure; see blg_last_error)";
    case BLG_E_NOMEM: return "out of memory";
    default: return "unknown status";
    }
}

int blg_create(int device, blg_ctx** out)
{
    if (!out) return BLG_E_INVALID;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || device < 0 || device >= ndev) return BLG_E_CUDA;
    if (cudaSetDevice(device) != cudaSuccess) return BLG_E_CUDA;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return BLG_E_CUDA;
    if (prop.major != 10) return BLG_E_CUDA;         // built for sm_100a only: any other device is "no usable device" (biolib_gpu.h)
    blg_ctx* c = new (std::nothrow) blg_ctx();
    if (!c) return BLG_E_NOMEM;
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    snprintf(c->name, sizeof(c->name), "%s", prop.name);
    if (cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess ||
        cudaHostAlloc(reinterpret_cast<void**>(&c->h_total), 64, cudaHostAllocMapped) != cudaSuccess ||
        cudaHostGetDevicePointer(reinterpret_cast<void**>(&c->d_total), c->h_total, 0) != cudaSuccess) {
        blg_destroy(c);
        return BLG_E_CUDA;
    }
    c->stream = c->own_stream;
    c->pool_dev.limit = (size_t)24 << 30;        // B200: 180 GB of HBM
    c->pool_dev.device = true;
    c->pool_pinned.limit = (size_t)2 << 30;
    bool ok = cudaStreamCreateWithFlags(&c->s_in, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&c->s_out, cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; ok && i < PIPE_MAX; ++i) ok = cudaEventCreateWithFlags(&c->ev_in[i], cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; ok && i < 2; ++i) ok = cudaEventCreateWithFlags(&c->ev_dl[i], cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; ok && i < PIPE_MAX; ++i) ok = cudaEventCreateWithFlags(&c->ev_pk[i], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) { blg_destroy(c); return BLG_E_CUDA; }
    *out = c;
    return BLG_OK;
}

int blg_destroy(blg_ctx* c)
{
    if (!c) return BLG_OK;
    cudaSetDevice(c->device);
    if (c->own_stream) cudaStreamSynchronize(c->own_stream);
    for (DevBuf* b : {&c->in, &c->offs, &c->status, &c->small, &c->packed, &c->brk, &c->fx_text, &c->fx_scratch, &c->sk4_status, &c->sk4_small, &c->o_value, &c->o_mm, &c->o_size, &c->o_pos, &c->o_hash,
                      &c->o_id, &c->rank, &c->o_lo, &c->o_hi, &c->o_position, &c->o_tlo, &c->o_thi, &c->o_tnib})
        if (b->p) cudaFree(b->p);
    for (auto& b : c->pool_dev.free_blocks) cudaFree(b.p);
    for (auto& b : c->pool_pinned.free_blocks) cudaFreeHost(b.p);
    if (c->h_total) cudaFreeHost(c->h_total);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    for (cudaEvent_t e : c->ev_in) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : c->ev_dl) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : c->ev_pk) if (e) cudaEventDestroy(e);
    if (c->s_in) cudaStreamDestroy(c->s_in);
    if (c->s_out) cudaStreamDestroy(c->s_out);
    delete c;
    return BLG_OK;
}

int blg_set_stream(blg_ctx* c, void* s)
{
    if (!c) return BLG_E_INVALID;
    c->stream = s ? static_cast<cudaStream_t>(s) : c->own_stream;
    return BLG_OK;
}

int blg_set_pipeline(blg_ctx* c, int chunk_mi, int first_div)
{
    if (!c || chunk_mi < 1 || chunk_mi > 1024 || first_div < 1 || first_div > 64) return BLG_E_INVALID;
    c->pipe_chunk_m = chunk_mi;
    c->pipe_first_div = first_div;
    return BLG_OK;
}

const char* blg_last_error(blg_ctx* c) { return c ? c->err.c_str() : "no context"; }

int blg_async_count(blg_ctx* c, uint64_t* count)
{
    if (!c || !count) return BLG_E_INVALID;
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaStreamSynchronize(c->stream));
    *count = c->h_total[0];
    c->stats.n_records = c->h_total[0];
    c->stats.exact_tiles = c->h_total[1] - c->sk4_exact_seen;     // of every launch since the last collected one
    c->sk4_exact_seen = c->h_total[1];
    add_kernel_time(c);
    if (c->async_capacity && *count > c->async_capacity) { c->err = "output capacity too small"; return BLG_E_CAPACITY; }
    return BLG_OK;
}

int blg_last_stats(blg_ctx* c, blg_stats* out)
{
    if (!c || !out) return BLG_E_INVALID;
    *out = c->stats;
    return BLG_OK;
}

int blg_device_info(blg_ctx* c, int* sm_count, char* name, size_t name_len)
{
    if (!c) return BLG_E_INVALID;
    if (sm_count) *sm_count = c->sm_count;
    if (name && name_len) snprintf(name, name_len, "%s", c->name);
    return BLG_OK;
}

int blg_host_alloc(blg_ctx* c, size_t bytes, void** out)
{
    if (!c || !out) return BLG_E_INVALID;
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaMallocHost(out, bytes ? bytes : 1));
    return BLG_OK;
}

int blg_host_free(blg_ctx* c, void* p)
{
    if (!c) return BLG_E_INVALID;
    if (p) CK(c, cudaFreeHost(p));
    return BLG_OK;
}

int blg_super_kmers_range(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, uint64_t win_begin,
                          uint64_t win_end, const blg_params* p, blg_sk_out* out)
{
    return sk_common(c, bases, n, offs, n_seq, win_begin, win_end, p, out, true);
}

int blg_super_kmers(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, const blg_params* p,
                    blg_sk_out* out)
{
    return sk_common(c, bases, n, offs, n_seq, 0, ~0ULL, p, out, true);
}

int blg_minimizers(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, const blg_params* p,
                   blg_sk_out* out)
{
    return sk_common(c, bases, n, offs, n_seq, 0, ~0ULL, p, out, false);
}

int blg_super_kmers_packed(blg_ctx* c, const blg_packed_seq* seq, const uint64_t* offs, size_t n_seq, uint64_t win_begin,
                           uint64_t win_end, const blg_params* p, blg_sk_out* out)
{
    if (!c || !seq || !p || !out || (seq->n_bases && !seq->codes) || (seq->n_breaks && (!seq->break_start || !seq->break_len))) return BLG_E_INVALID;
    if (p->flags & (BLG_F_DEVICE_INPUT | BLG_F_EMIT_ID)) { c->err = "packed input: host pointers only, no BLG_F_EMIT_ID"; return BLG_E_INVALID; }
    c->err.clear();
    CK(c, cudaSetDevice(c->device));
    const size_t n = seq->n_bases, words = (n + 15) / 16;
    int rc;
    if ((rc = ensure(c, c->packed, words * 4 + 16)) != BLG_OK || (rc = ensure(c, c->in, words * 16 + 64)) != BLG_OK) return rc;
    if ((rc = ensure(c, c->brk, seq->n_breaks * 12 + 16)) != BLG_OK) return rc;
    c->stats = blg_stats{};
    unsigned long long* d_bs = static_cast<unsigned long long*>(c->brk.p);
    unsigned int* d_bl = reinterpret_cast<unsigned int*>(d_bs + seq->n_breaks);
    // inputs that take the chunked host path (same test as sk_pipelined's) are uploaded and unpacked there, piece by piece, overlapped with
    // the kernels and the record download; everything else is staged here in one go
    uint64_t hi = n >= (size_t)p->k ? n - p->k + 1 : 0;
    if (win_end < hi) hi = win_end;
    const uint64_t lo = win_begin < hi ? win_begin : hi;
    const bool lazy = !(p->flags & BLG_F_DEVICE_OUTPUT) && hi - lo >= (8u << 20);
    cudaStream_t up = lazy ? c->s_in : c->stream;
    if (seq->n_breaks) {
        CK(c, cudaMemcpyAsync(d_bs, seq->break_start, seq->n_breaks * 8, cudaMemcpyHostToDevice, up));
        CK(c, cudaMemcpyAsync(d_bl, seq->break_len, seq->n_breaks * 4, cudaMemcpyHostToDevice, up));
    }
    c->stats.bytes_h2d = words * 4 + seq->n_breaks * 12;
    if (!lazy) {
        if (words) CK(c, cudaMemcpyAsync(c->packed.p, seq->codes, words * 4, cudaMemcpyHostToDevice, c->stream));
        CK(c, unpack_launch(c->packed.p, n, d_bs, d_bl, seq->n_breaks, static_cast<unsigned char*>(c->in.p), c->stream));
        c->stats.kernel_launches = seq->n_breaks ? 2 : 1;
    }
    c->staged = static_cast<const unsigned char*>(c->in.p);
    c->packed_lazy = lazy ? seq : nullptr;
    c->lazy_breaks = seq->n_breaks;
    rc = sk_common(c, reinterpret_cast<const char*>(c->in.p), n, offs, n_seq, win_begin, win_end, p, out, true);
    c->staged = nullptr;
    c->packed_lazy = nullptr;
    return rc;
}

int blg_super_kmer_payloads(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, const blg_params* p,
                            uint32_t n_buckets, blg_payload_out* out)
{
    if (!c || !p || !out) return BLG_E_INVALID;
    if (n_buckets == 0 || n_buckets > 4096) { c->err = "need 1 <= n_buckets <= 4096"; return BLG_E_INVALID; }
    if (p->flags & (BLG_F_EMIT_ID | BLG_F_EMIT_HASH | BLG_F_POS32 | BLG_F_NO_SIZE)) { c->err = "payloads: only BLG_F_DEVICE_INPUT / BLG_F_DEVICE_OUTPUT apply"; return BLG_E_INVALID; }
    const bool dev_out = (p->flags & BLG_F_DEVICE_OUTPUT) != 0;
    *out = blg_payload_out{};
    out->n_buckets = n_buckets;
    // 1. the records, device-resident (library-owned arrays)
    blg_params q = *p;
    q.flags = (p->flags & BLG_F_DEVICE_INPUT) | BLG_F_DEVICE_OUTPUT;
    blg_sk_out rec{};
    int rc = sk_common(c, bases, n, offs, n_seq, 0, ~0ULL, &q, &rec, true);
    if (rc != BLG_OK) return rc;
    Timer tm;
    const blg_stats sk_stats = c->stats;
    const uint64_t cnt = rec.count;
    const int k = p->k, w = p->k - p->m + 1;
    const unsigned char* d_bases = (p->flags & BLG_F_DEVICE_INPUT) ? reinterpret_cast<const unsigned char*>(bases) : static_cast<const unsigned char*>(c->in.p);
    Owner* dv = new (std::nothrow) Owner{1, {}, {}, nullptr};
    Owner* rec_owner = static_cast<Owner*>(rec.owner);
    auto fail = [&](int code) { free_owner(dv); free_owner(rec_owner); return code; };
    if (!dv) return fail(BLG_E_NOMEM);
    const uint64_t max_bases = cnt * (uint64_t)(w + k - 1);
    const size_t max_words = (size_t)(max_bases / 32 + 2);
    unsigned long long *d_bo = nullptr, *d_v = nullptr, *d_mm = nullptr, *d_ps = nullptr, *d_words = nullptr;
    unsigned char *d_sz = nullptr, *d_mp = nullptr;
    if ((rc = alloc_owned(c, true, (size_t)n_buckets + 1, &d_bo, dv, 0)) != BLG_OK || (rc = alloc_owned(c, true, cnt, &d_v, dv, 1)) != BLG_OK ||
        (rc = alloc_owned(c, true, cnt, &d_mm, dv, 2)) != BLG_OK || (rc = alloc_owned(c, true, cnt, &d_sz, dv, 3)) != BLG_OK ||
        (rc = alloc_owned(c, true, cnt, &d_mp, dv, 4)) != BLG_OK || (rc = alloc_owned(c, true, cnt, &d_ps, dv, 5)) != BLG_OK ||
        (rc = alloc_owned(c, true, max_words, &d_words, dv, 6)) != BLG_OK) return fail(rc);
    if ((rc = ensure(c, c->rank, payload_scratch_bytes(cnt, n_buckets) + 64)) != BLG_OK) return fail(rc);
    if (cudaMemsetAsync(d_words, 0, max_words * 8, c->stream) != cudaSuccess) return fail(BLG_E_CUDA);
    unsigned long long* d_tot = reinterpret_cast<unsigned long long*>(static_cast<char*>(c->rank.p) + payload_scratch_bytes(cnt, n_buckets) / 8 * 8);
    int launches = 0;
    cudaEventRecord(c->ev0, c->stream);
    cudaError_t e = payload_launch(d_bases, (unsigned long long)n, reinterpret_cast<unsigned long long*>(rec.value), reinterpret_cast<unsigned long long*>(rec.mm_start), rec.size,
                                   rec.mm_pos, cnt, k, (unsigned long long)(uint32_t)p->seed ^ 8ULL, n_buckets, c->rank.p, d_v, d_mm, d_sz, d_mp, d_ps,
                                   d_bo, d_words, d_tot, c->stream, &launches);
    cudaEventRecord(c->ev1, c->stream);
    unsigned long long tot[2] = {0, 0};
    if (e != cudaSuccess || cudaMemcpyAsync(tot, d_tot, 16, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess) { c->err = "payload kernels failed"; return fail(BLG_E_CUDA); }
    free_owner(rec_owner);
    rec_owner = nullptr;
    c->stats = sk_stats;
    c->stats.kernel_launches += launches;
    add_kernel_time(c);
    out->count = cnt;
    out->total_bases = tot[0];
    out->n_words = (size_t)(tot[0] * 2 / 64 + 1);                 // bit::packed::vector::resize_data: size * bitwidth / 64 + 1
    if (dev_out) {
        out->bucket_offsets = reinterpret_cast<uint64_t*>(d_bo); out->value = reinterpret_cast<uint64_t*>(d_v);
        out->mm_start = reinterpret_cast<uint64_t*>(d_mm); out->size = d_sz; out->mm_pos = d_mp;
        out->payload_start = reinterpret_cast<uint64_t*>(d_ps); out->words = reinterpret_cast<uint64_t*>(d_words);
        out->owner = dv;
        c->stats.total_ms += tm.ms();
        return BLG_OK;
    }
    Owner* hv = new (std::nothrow) Owner{0, {}, {}, nullptr};
    if (!hv) return fail(BLG_E_NOMEM);
    auto fail2 = [&](int code) { free_owner(hv); return fail(code); };
    if ((rc = alloc_owned(c, false, (size_t)n_buckets + 1, &out->bucket_offsets, hv, 0)) != BLG_OK || (rc = alloc_owned(c, false, cnt, &out->value, hv, 1)) != BLG_OK ||
        (rc = alloc_owned(c, false, cnt, &out->mm_start, hv, 2)) != BLG_OK || (rc = alloc_owned(c, false, cnt, &out->size, hv, 3)) != BLG_OK ||
        (rc = alloc_owned(c, false, cnt, &out->mm_pos, hv, 4)) != BLG_OK || (rc = alloc_owned(c, false, cnt, &out->payload_start, hv, 5)) != BLG_OK ||
        (rc = alloc_owned(c, false, out->n_words, &out->words, hv, 6)) != BLG_OK) return fail2(rc);
    bool ok = cudaMemcpyAsync(out->bucket_offsets, d_bo, ((size_t)n_buckets + 1) * 8, cudaMemcpyDeviceToHost, c->stream) == cudaSuccess;
    if (cnt) {
        ok = ok && cudaMemcpyAsync(out->value, d_v, cnt * 8, cudaMemcpyDeviceToHost, c->stream) == cudaSuccess;
        ok = ok && cudaMemcpyAsync(out->mm_start, d_mm, cnt * 8, cudaMemcpyDeviceToHost, c->stream) == cudaSuccess;
        ok = ok && cudaMemcpyAsync(out->size, d_sz, cnt, cudaMemcpyDeviceToHost, c->stream) == cudaSuccess;
        ok = ok && cudaMemcpyAsync(out->mm_pos, d_mp, cnt, cudaMemcpyDeviceToHost, c->stream) == cudaSuccess;
        ok = ok && cudaMemcpyAsync(out->payload_start, d_ps, cnt * 8, cudaMemcpyDeviceToHost, c->stream) == cudaSuccess;
    }
    ok = ok && cudaMemcpyAsync(out->words, d_words, out->n_words * 8, cudaMemcpyDeviceToHost, c->stream) == cudaSuccess;
    ok = ok && cudaStreamSynchronize(c->stream) == cudaSuccess;
    if (!ok) { c->err = "payload download failed"; return fail2(BLG_E_CUDA); }
    c->stats.bytes_d2h += ((size_t)n_buckets + 1) * 8 + cnt * 26 + out->n_words * 8;
    out->owner = hv;
    free_owner(dv);
    c->stats.total_ms += tm.ms();
    return BLG_OK;
}

void blg_free_payload_out(blg_ctx* c, blg_payload_out* out)
{
    if (!out) return;
    if (c) cudaSetDevice(c->device);
    if (out->owner) free_owner(static_cast<Owner*>(out->owner));
    *out = blg_payload_out{};
}

int blg_parse_fastx(blg_ctx* c, const char* text, size_t n_text, int format, uint32_t flags, blg_fastx_out* out)
{
    if (!c || !out || (n_text && !text) || format < -1 || format > 1) return BLG_E_INVALID;
    c->err.clear();
    c->stats = blg_stats{};
    Timer tm;
    CK(c, cudaSetDevice(c->device));
    out->n_bases = out->n_seq = 0; out->bases = nullptr; out->seq_offsets = nullptr; out->owner = nullptr;
    if (n_text == 0) return BLG_OK;
    const bool dev_in = (flags & BLG_F_DEVICE_INPUT) != 0, dev_out = (flags & BLG_F_DEVICE_OUTPUT) != 0;
    int rc;
    const unsigned char* d_text = reinterpret_cast<const unsigned char*>(text);
    if (!dev_in) {
        if ((rc = ensure(c, c->fx_text, n_text + 64)) != BLG_OK) return rc;
        CK(c, cudaMemcpyAsync(c->fx_text.p, text, n_text, cudaMemcpyHostToDevice, c->stream));
        c->stats.bytes_h2d += n_text;
        d_text = static_cast<const unsigned char*>(c->fx_text.p);
    } else if (reinterpret_cast<uintptr_t>(text) & 15u) { c->err = "device text must be 16-byte aligned"; return BLG_E_INVALID; }
    if (format < 0) {
        char first = 0;
        if (dev_in) CK(c, cudaMemcpy(&first, text, 1, cudaMemcpyDeviceToHost)); else first = text[0];
        if (first == '>') format = 0; else if (first == '@') format = 1;
        else { c->err = "text starts with neither '>' nor '@'"; return BLG_E_INVALID; }
    }
    if ((rc = ensure(c, c->fx_scratch, fastx_scratch_bytes(n_text))) != BLG_OK) return rc;
    // totals and the first record's offset arrive in the context's mapped pinned words [4..6]: no copy, one synchronise each
    volatile unsigned long long* h_fx = c->h_total + 4;
    h_fx[0] = h_fx[1] = 0; h_fx[2] = 1;
    int launches = 0;
    cudaError_t e = fastx_count_launch(d_text, n_text, format, c->fx_scratch.p, c->d_total + 4, c->stream, &launches);
    if (e != cudaSuccess || cudaStreamSynchronize(c->stream) != cudaSuccess) { c->err = "fastx: counting passes failed"; return BLG_E_CUDA; }
    const unsigned long long tot[2] = {h_fx[0], h_fx[1]};
    if (tot[1] == 0) { c->err = "text does not start with a record header"; return BLG_E_INVALID; }
    Owner* ow = new (std::nothrow) Owner{1, {}, {}, nullptr};
    if (!ow) return BLG_E_NOMEM;
    unsigned char* d_bases = nullptr;
    unsigned long long* d_offs = nullptr;
    if ((rc = alloc_owned(c, true, (size_t)tot[0] + 64, &d_bases, ow, 0)) != BLG_OK || (rc = alloc_owned(c, true, (size_t)tot[1] + 1, &d_offs, ow, 1)) != BLG_OK) { free_owner(ow); return rc; }
    e = fastx_emit_launch(d_text, n_text, format, c->fx_scratch.p, d_bases, d_offs, c->d_total + 6, c->stream, &launches);
    if (e != cudaSuccess) { free_owner(ow); c->err = cudaGetErrorString(e); return BLG_E_CUDA; }
    c->stats.kernel_launches += launches;
    if (dev_out && cudaStreamSynchronize(c->stream) != cudaSuccess) { free_owner(ow); c->err = "fastx: scatter pass failed"; return BLG_E_CUDA; }
    out->n_bases = tot[0];
    out->n_seq = tot[1];
    c->stats.n_bases = tot[0];
    c->stats.n_records = tot[1];
    if (dev_out) {
        out->bases = reinterpret_cast<char*>(d_bases);
        out->seq_offsets = reinterpret_cast<uint64_t*>(d_offs);
        out->owner = ow;
    } else {
        Owner* hw = new (std::nothrow) Owner{0, {}, {}, nullptr};
        if (!hw) { free_owner(ow); return BLG_E_NOMEM; }
        char* hb = nullptr;
        uint64_t* ho = nullptr;
        if ((rc = alloc_owned(c, false, (size_t)tot[0] + 1, &hb, hw, 0)) != BLG_OK || (rc = alloc_owned(c, false, (size_t)tot[1] + 1, &ho, hw, 1)) != BLG_OK) { free_owner(ow); free_owner(hw); return rc; }
        cudaError_t e1 = cudaMemcpyAsync(hb, d_bases, tot[0], cudaMemcpyDeviceToHost, c->stream);
        cudaError_t e2 = cudaMemcpyAsync(ho, d_offs, (tot[1] + 1) * 8, cudaMemcpyDeviceToHost, c->stream);
        cudaError_t e3 = cudaStreamSynchronize(c->stream);
        free_owner(ow);
        if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess) { free_owner(hw); c->err = "fastx: download failed"; return BLG_E_CUDA; }
        c->stats.bytes_d2h += tot[0] + (tot[1] + 1) * 8;
        out->bases = hb; out->seq_offsets = ho; out->owner = hw;
    }
    if (h_fx[2] != 0) {                      // bytes of a sequence line before the first header
        blg_free_fastx_out(c, out);
        c->err = "text does not start with a record header";
        return BLG_E_INVALID;
    }
    c->stats.total_ms = tm.ms();
    return BLG_OK;
}

void blg_free_fastx_out(blg_ctx* c, blg_fastx_out* out)
{
    if (!out) return;
    if (c) cudaSetDevice(c->device);
    if (out->owner) free_owner(static_cast<Owner*>(out->owner));
    out->bases = nullptr; out->seq_offsets = nullptr; out->owner = nullptr; out->n_bases = out->n_seq = 0;
}

namespace {
// Host input, host output, 8 Mi k-mer starts or more: pieces of pipe_chunk_m Mi starts whose upload (s_in), kernel (stream) and
// record download (s_out) overlap. Two device piece buffers alternate, so the device holds two pieces of output rather than
// one record slot per k-mer start of the whole input. `est` sizes library-owned host arrays; when a piece does not fit what
// is left of them (or of the caller's capacity) *done_flag stays 0 and the caller takes the plain path, which knows the count.
int kmers_pipelined(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, const KmerArgs& proto, uint64_t est,
                    blg_kmer_out* out, int* done_flag, bool runs)
{
    *done_flag = 0;
    std::vector<uint64_t> run_first, run_start;     // BLG_F_RUNS: collected piece by piece
    const int k = proto.k;
    const bool wide = k > 32;
    const uint64_t nwin = n - (size_t)k + 1;
    if (nwin < (8u << 20) || c->staged) return BLG_OK;
    int rc;
    if ((rc = check_offsets_host(offs, n_seq, n)) != BLG_OK) { c->err = "seq_offsets must start at 0, end at n_bases and be non-decreasing"; return rc; }
    const uint64_t cap = out->capacity > 0 ? out->capacity : est;
    int nch = (int)(nwin / ((uint64_t)c->pipe_chunk_m << 20));
    nch = nch < 2 ? 2 : (nch > PIPE_MAX ? PIPE_MAX : nch);
    uint64_t cut[PIPE_MAX + 1];
    size_t up[PIPE_MAX + 1];
    const uint64_t first = nwin / nch / (uint64_t)c->pipe_first_div;
    uint64_t piece = 0;
    for (int i = 0; i <= nch; ++i) {
        cut[i] = i == nch ? nwin : i == 0 ? 0 : (first + (nwin - first) * (uint64_t)(i - 1) / (nch - 1)) / 16 * 16;
        if (i && cut[i] - cut[i - 1] > piece) piece = cut[i] - cut[i - 1];
    }
    if ((rc = ensure(c, c->in, n + 64)) != BLG_OK) return rc;
    if ((rc = ensure(c, c->o_lo, 2 * piece * 8)) != BLG_OK || (rc = ensure(c, c->o_position, 2 * piece * 8)) != BLG_OK) return rc;
    if (wide && (rc = ensure(c, c->o_hi, 2 * piece * 8)) != BLG_OK) return rc;
    const unsigned long long* d_offs = nullptr;
    if (offs) {
        const size_t ob = (n_seq + 1) * sizeof(uint64_t);
        if ((rc = ensure(c, c->offs, ob)) != BLG_OK) return rc;
        CK(c, cudaMemcpyAsync(c->offs.p, offs, ob, cudaMemcpyHostToDevice, c->stream));
        c->stats.bytes_h2d += ob;
        d_offs = static_cast<const unsigned long long*>(c->offs.p);
    }
    Owner* ow = nullptr;
    struct Guard {
        blg_ctx* c; Owner** ow; bool armed;
        ~Guard() { if (armed) { cudaStreamSynchronize(c->s_in); cudaStreamSynchronize(c->s_out); cudaStreamSynchronize(c->stream); free_owner(*ow); *ow = nullptr; } }
    } guard{c, &ow, true};
    blg_kmer_out h = *out;
    if (out->capacity == 0) {
        ow = new (std::nothrow) Owner{0, {}, {}, nullptr};
        if (!ow) return BLG_E_NOMEM;
        h.value_hi = nullptr; h.position = nullptr;
        if ((rc = alloc_owned(c, false, cap, &h.value_lo, ow, 0)) != BLG_OK || (!runs && (rc = alloc_owned(c, false, cap, &h.position, ow, 2)) != BLG_OK)) return rc;
        if (wide && (rc = alloc_owned(c, false, cap, &h.value_hi, ow, 1)) != BLG_OK) return rc;
    }
    if (runs) h.position = nullptr;                 // never downloaded
    unsigned char* d_in = static_cast<unsigned char*>(c->in.p);
    const size_t halo = (size_t)k + 64;
    size_t sent = 0;
    for (int i = 0; i < nch; ++i) {
        size_t end = i == nch - 1 ? n : (size_t)cut[i + 1] + halo;
        end = end > n ? n : end;
        if (i < nch - 1) end = end / 16 * 16;
        if (end > sent) CK(c, cudaMemcpyAsync(d_in + sent, bases + sent, end - sent, cudaMemcpyHostToDevice, c->s_in));
        CK(c, cudaEventRecord(c->ev_in[i], c->s_in));
        sent = end > sent ? end : sent;
        up[i] = sent;
    }
    c->stats.bytes_h2d += n;
    uint64_t done = 0;
    bool overflow = false;
    int used[2] = {0, 0};
    for (int i = 0, b = 0; i < nch; ++i) {
        if (cut[i + 1] <= cut[i]) continue;
        CK(c, cudaStreamWaitEvent(c->stream, c->ev_in[i], 0));
        if (used[b]) CK(c, cudaStreamWaitEvent(c->stream, c->ev_dl[b], 0));    // the piece before last has left this buffer
        KmerArgs a = proto;
        a.bases = d_in; a.buf_lo = 0; a.buf_hi = (long long)up[i];
        a.seq_offsets = d_offs; a.n_seq = (long long)n_seq;
        a.win_lo = (long long)cut[i]; a.win_hi = (long long)cut[i + 1];
        a.vec_ok = (reinterpret_cast<uintptr_t>(d_in) & 15u) == 0;
        a.value_lo = static_cast<unsigned long long*>(c->o_lo.p) + (size_t)b * piece;
        a.value_hi = wide ? static_cast<unsigned long long*>(c->o_hi.p) + (size_t)b * piece : nullptr;
        a.position = static_cast<unsigned long long*>(c->o_position.p) + (size_t)b * piece;
        a.capacity = piece;
        uint64_t cnt = 0;
        if ((rc = run_kmer_once(c, a, &cnt)) != BLG_OK) return rc;      // returns after the kernel: the count is known
        if (cnt > cap - done) { overflow = true; break; }
        if (cnt) {
            CK(c, cudaMemcpyAsync(h.value_lo + done, a.value_lo, cnt * 8, cudaMemcpyDeviceToHost, c->s_out));
            if (wide) CK(c, cudaMemcpyAsync(h.value_hi + done, a.value_hi, cnt * 8, cudaMemcpyDeviceToHost, c->s_out));
            if (h.position) CK(c, cudaMemcpyAsync(h.position + done, a.position, cnt * 8, cudaMemcpyDeviceToHost, c->s_out));
            c->stats.bytes_d2h += cnt * ((wide ? 16 : 8) + (h.position ? 8 : 0));
            CK(c, cudaEventRecord(c->ev_dl[b], c->s_out));
            // the positions stay on the device: their run list (a few entries per break / sequence) is all that is downloaded
            if (runs && (rc = append_kmer_runs(c, a.position, cnt, done, run_first, run_start)) != BLG_OK) return rc;
            used[b] = 1;
            b ^= 1;
        }
        done += cnt;
    }
    CK(c, cudaStreamSynchronize(c->s_out));
    CK(c, cudaStreamSynchronize(c->s_in));
    if (overflow) return BLG_OK;       // the guard gives library arrays back; the plain path reports the exact count
    if (runs) {
        h.run_capacity = out->run_capacity; h.run_first_record = out->run_first_record; h.run_start = out->run_start;
        if ((rc = deliver_kmer_runs(c, run_first, run_start, &h, out->capacity == 0 ? ow : nullptr)) != BLG_OK) {
            if (rc == BLG_E_CAPACITY) { out->n_runs = h.n_runs; out->count = done; }
            return rc;
        }
        if (out->capacity > 0) out->n_runs = h.n_runs;
    }
    guard.armed = false;
    if (out->capacity == 0) { *out = h; out->owner = ow; out->capacity = 0; }
    out->count = done;
    c->stats.n_records = done;
    *done_flag = 1;
    return BLG_OK;
}

// sync_s < 0: every valid k-mer (blg_kmers); else the closed-syncmer filter with m-mers of p->m bases (blg_syncmers);
// threshold != nullptr: the hash-threshold filter (blg_hash_sampled_kmers)
int kmers_common(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, const blg_params* p,
                 int sync_s, int sync_e, blg_kmer_out* out, const uint64_t* threshold = nullptr, bool want_pos = true)
{
    // want_pos = false (internal callers that only sort the values): library-owned device output without the position array
    if (!c || !p || !out) return BLG_E_INVALID;
    if (p->k == 0 || p->k > 64) { c->err = "need 1 <= k <= 64"; return BLG_E_INVALID; }
    if (sync_s >= 0 && (p->k > 32 || p->m == 0 || p->m > p->k)) { c->err = "syncmers need 1 <= m <= k <= 32"; return BLG_E_INVALID; }
    if (threshold && p->k > 32) { c->err = "hash sampling needs k <= 32 (hash64 of a 64-bit value)"; return BLG_E_INVALID; }
    const bool wide = p->k > 32;
    if (out->capacity > 0 && (!out->value_lo || (wide && !out->value_hi))) { c->err = "value arrays missing"; return BLG_E_INVALID; }
    const bool runs = want_pos && (p->flags & BLG_F_RUNS) != 0;    // positions stay on the device; their run list is downloaded
    if (runs && (sync_s >= 0 || threshold || (p->flags & BLG_F_DEVICE_OUTPUT))) {
        c->err = "BLG_F_RUNS: host output of blg_kmers only (sampled k-mers are not contiguous)";
        return BLG_E_INVALID;
    }
    Timer tm;
    c->stats = blg_stats{};
    c->stats.n_bases = n;
    c->err.clear();
    CK(c, cudaSetDevice(c->device));
    const bool dev_out = (p->flags & BLG_F_DEVICE_OUTPUT) != 0;
    const int k = p->k;
    const uint64_t nwin = n >= (size_t)k ? n - k + 1 : 0;
    out->count = 0;
    if (runs) out->n_runs = 0;
    if (nwin == 0) {
        if (out->capacity == 0) {
            out->value_lo = out->value_hi = out->position = nullptr; out->owner = nullptr;
            if (runs) out->run_first_record = out->run_start = nullptr;
        }
        c->stats.total_ms = tm.ms();
        return BLG_OK;
    }
    int rc;
    KmerArgs a{};
    a.k = k; a.canonical = p->canonical ? 1 : 0;
    if (sync_s >= 0) { a.sync_on = 1; a.sync_m = p->m; a.sync_s = sync_s; a.sync_e = sync_e; }
    if (threshold) { a.sync_on = 2; a.samp_s8 = (unsigned long long)(uint32_t)p->seed ^ 8ULL; a.samp_threshold = *threshold; }
    if (!(p->flags & (BLG_F_DEVICE_INPUT | BLG_F_DEVICE_OUTPUT))) {
        // expected number of kept k-mers (+25 % and a constant): every start; two window offsets out of k - m + 1; the sampled fraction
        uint64_t est = nwin;
        if (sync_s >= 0) est = (uint64_t)((double)nwin * (sync_s == sync_e ? 1.0 : 2.0) / (double)(k - p->m + 1) * 1.25) + 65536;
        if (threshold) est = (uint64_t)((double)nwin * ((double)*threshold / 18446744073709551616.0) * 1.25) + 65536;
        if (est > nwin) est = nwin;
        int produced = 0;
        if ((rc = kmers_pipelined(c, bases, n, offs, n_seq, a, est, out, &produced, runs)) != BLG_OK) return rc;
        if (produced) { c->stats.total_ms = tm.ms(); return BLG_OK; }
    }
    const unsigned char* d_bases = nullptr;
    const unsigned long long* d_offs = nullptr;
    if ((rc = place_input(c, bases, n, offs, n_seq, p->flags, &d_bases, &d_offs)) != BLG_OK) return rc;
    a.bases = d_bases; a.buf_lo = 0; a.buf_hi = (long long)n;
    a.seq_offsets = d_offs; a.n_seq = (long long)n_seq;
    a.win_lo = 0; a.win_hi = (long long)nwin;
    a.vec_ok = (reinterpret_cast<uintptr_t>(d_bases) & 15u) == 0;
    uint64_t total = 0;
    if (dev_out && out->capacity > 0) {
        a.value_lo = reinterpret_cast<unsigned long long*>(out->value_lo);
        a.value_hi = wide ? reinterpret_cast<unsigned long long*>(out->value_hi) : nullptr;
        a.position = reinterpret_cast<unsigned long long*>(out->position);
        a.capacity = out->capacity;
        if ((rc = run_kmer_once(c, a, &total)) != BLG_OK) return rc;
        out->count = total;
        c->stats.n_records = total;
        c->stats.total_ms = tm.ms();
        if (total > out->capacity) { c->err = "output capacity too small"; return BLG_E_CAPACITY; }
        return BLG_OK;
    }
    // at most one record per k-mer start: size the device arrays for that, no retry needed
    uint64_t cap = nwin;
    if (!dev_out && out->capacity > 0 && out->capacity < cap) cap = out->capacity;
    Owner* ow = nullptr;
    if (dev_out) {
        ow = new (std::nothrow) Owner{1, {}, {}, nullptr};
        if (!ow) return BLG_E_NOMEM;
        unsigned long long *lo = nullptr, *hi = nullptr, *pos = nullptr;
        if ((rc = alloc_owned(c, true, cap, &lo, ow, 0)) != BLG_OK || (want_pos && (rc = alloc_owned(c, true, cap, &pos, ow, 2)) != BLG_OK)) { free_owner(ow); return rc; }
        if (wide && (rc = alloc_owned(c, true, cap, &hi, ow, 1)) != BLG_OK) { free_owner(ow); return rc; }
        a.value_lo = lo; a.value_hi = hi; a.position = pos;
    } else {
        if ((rc = ensure(c, c->o_lo, cap * 8)) != BLG_OK || (rc = ensure(c, c->o_position, cap * 8)) != BLG_OK) return rc;
        if (wide && (rc = ensure(c, c->o_hi, cap * 8)) != BLG_OK) return rc;
        a.value_lo = static_cast<unsigned long long*>(c->o_lo.p);
        a.value_hi = wide ? static_cast<unsigned long long*>(c->o_hi.p) : nullptr;
        a.position = static_cast<unsigned long long*>(c->o_position.p);
    }
    a.capacity = cap;
    if ((rc = run_kmer_once(c, a, &total)) != BLG_OK) { free_owner(ow); return rc; }
    out->count = total;
    c->stats.n_records = total;
    if (dev_out) {
        out->value_lo = reinterpret_cast<uint64_t*>(a.value_lo);
        out->value_hi = reinterpret_cast<uint64_t*>(a.value_hi);
        out->position = reinterpret_cast<uint64_t*>(a.position);
        out->owner = ow;
        c->stats.total_ms = tm.ms();
        return BLG_OK;
    }
    if (out->capacity > 0) {
        if (total > out->capacity) { c->stats.total_ms = tm.ms(); c->err = "output capacity too small"; return BLG_E_CAPACITY; }
    } else {
        ow = new (std::nothrow) Owner{0, {}, {}, nullptr};
        if (!ow) return BLG_E_NOMEM;
        out->value_hi = nullptr; out->position = nullptr;
        if ((rc = alloc_owned(c, false, total, &out->value_lo, ow, 0)) != BLG_OK || (!runs && (rc = alloc_owned(c, false, total, &out->position, ow, 2)) != BLG_OK)) { free_owner(ow); return rc; }
        if (wide && (rc = alloc_owned(c, false, total, &out->value_hi, ow, 1)) != BLG_OK) { free_owner(ow); return rc; }
        out->owner = ow;
    }
    if (total) {
        const bool dl_pos = !runs && out->position != nullptr;
        CK(c, cudaMemcpyAsync(out->value_lo, a.value_lo, total * 8, cudaMemcpyDeviceToHost, c->stream));
        if (wide) CK(c, cudaMemcpyAsync(out->value_hi, a.value_hi, total * 8, cudaMemcpyDeviceToHost, c->stream));
        if (dl_pos) CK(c, cudaMemcpyAsync(out->position, a.position, total * 8, cudaMemcpyDeviceToHost, c->stream));
        c->stats.bytes_d2h += total * ((wide ? 16 : 8) + (dl_pos ? 8 : 0));
        CK(c, cudaStreamSynchronize(c->stream));
    }
    if (runs) {
        std::vector<uint64_t> run_first, run_start;
        if ((rc = append_kmer_runs(c, a.position, total, 0, run_first, run_start)) != BLG_OK ||
            (rc = deliver_kmer_runs(c, run_first, run_start, out, out->capacity == 0 ? ow : nullptr)) != BLG_OK) {
            if (out->capacity == 0) { free_owner(ow); out->owner = nullptr; out->value_lo = out->value_hi = nullptr; out->run_first_record = out->run_start = nullptr; }
            return rc;
        }
    }
    c->stats.total_ms = tm.ms();
    return BLG_OK;
}
}  // namespace

int blg_kmers(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, const blg_params* p,
              blg_kmer_out* out)
{
    return kmers_common(c, bases, n, offs, n_seq, p, -1, -1, out);
}

int blg_syncmers(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, const blg_params* p,
                 uint16_t start_offset, uint16_t end_offset, blg_kmer_out* out)
{
    return kmers_common(c, bases, n, offs, n_seq, p, (int)start_offset, (int)end_offset, out);
}

int blg_hash_sampled_kmers(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, const blg_params* p,
                           uint64_t threshold, blg_kmer_out* out)
{
    return kmers_common(c, bases, n, offs, n_seq, p, -1, -1, out, &threshold);
}

namespace {
// distinct k-mer values of one input, sorted, in a fresh device array (*d_out, owned by *ow slot 0); *count = their number
int sorted_unique_device(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, const blg_params* p, Owner** ow_out,
                         unsigned long long** d_out, uint64_t* count, blg_stats* acc)
{
    *ow_out = nullptr; *d_out = nullptr; *count = 0;
    if (p->k == 0 || p->k > 32) { c->err = "sorted / unique / Jaccard: need 1 <= k <= 32 (64-bit k-mer values)"; return BLG_E_INVALID; }
    blg_params q = *p;
    q.flags = (p->flags & BLG_F_DEVICE_INPUT) | BLG_F_DEVICE_OUTPUT;
    blg_kmer_out km{};
    int rc = kmers_common(c, bases, n, offs, n_seq, &q, -1, -1, &km, nullptr, (p->flags & BLG_F_REF_DROP_LAST) != 0);
    if (rc != BLG_OK) return rc;
    acc->bytes_h2d += c->stats.bytes_h2d; acc->kernel_launches += c->stats.kernel_launches; acc->kernel_ms += c->stats.kernel_ms; acc->n_bases += n;
    Owner* kow = static_cast<Owner*>(km.owner);
    uint64_t cnt = km.count;
    Owner* ow = new (std::nothrow) Owner{1, {}, {}, nullptr};
    auto fail = [&](int code) { free_owner(ow); free_owner(kow); return code; };
    if (!ow) return fail(BLG_E_NOMEM);
    if (cnt == 0) { free_owner(kow); *ow_out = ow; return BLG_OK; }
    unsigned long long *d_tmp = nullptr, *d_flags = nullptr, *d_uniq = nullptr;
    if ((rc = alloc_owned(c, true, cnt, &d_uniq, ow, 0)) != BLG_OK || (rc = alloc_owned(c, true, cnt, &d_tmp, ow, 1)) != BLG_OK) return fail(rc);
    if ((p->flags & BLG_F_REF_DROP_LAST) && (rc = alloc_owned(c, true, cnt, &d_flags, ow, 2)) != BLG_OK) return fail(rc);   // the K1 filter's keep flags
    if ((rc = ensure(c, c->rank, sort_scratch_bytes(cnt) + 64)) != BLG_OK) return fail(rc);
    unsigned long long* d_tot = reinterpret_cast<unsigned long long*>(static_cast<char*>(c->rank.p) + sort_scratch_bytes(cnt) / 8 * 8);
    unsigned long long* vals = reinterpret_cast<unsigned long long*>(km.value_lo);
    int launches = 0;
    cudaEventRecord(c->ev0, c->stream);
    if (p->flags & BLG_F_REF_DROP_LAST) {
        const unsigned long long* d_offs = (p->flags & BLG_F_DEVICE_INPUT) ? reinterpret_cast<const unsigned long long*>(offs)
                                                                            : static_cast<const unsigned long long*>(offs ? c->offs.p : nullptr);
        if (drop_last_kmers(vals, reinterpret_cast<unsigned long long*>(km.position), cnt, p->k, d_offs, n_seq, n, d_flags, c->rank.p, d_tmp, d_tot,
                            c->stream, &launches) != cudaSuccess) return fail(BLG_E_CUDA);
        unsigned long long kept = 0;
        if (cudaMemcpyAsync(&kept, d_tot, 8, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess || cudaStreamSynchronize(c->stream) != cudaSuccess) return fail(BLG_E_CUDA);
        cnt = kept;
    }
    unsigned long long uniq = 0;
    if (cnt) {
        if (radix_sort_u64(vals, d_tmp, cnt, 2 * p->k, c->rank.p, c->stream, &launches) != cudaSuccess ||
            unique_sorted_u64(vals, cnt, d_flags, c->rank.p, d_uniq, d_tot, c->stream, &launches) != cudaSuccess) return fail(BLG_E_CUDA);
        cudaEventRecord(c->ev1, c->stream);
        if (cudaMemcpyAsync(&uniq, d_tot, 8, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess || cudaStreamSynchronize(c->stream) != cudaSuccess) {
            c->err = "sort / unique failed"; return fail(BLG_E_CUDA);
        }
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) == cudaSuccess) acc->kernel_ms += ms;
    }
    acc->kernel_launches += launches;
    free_owner(kow);
    // keep only the result array
    for (int i = 1; i <= 2; ++i) {
        if (!ow->ptrs[i]) continue;
        if (!c->pool_dev.give(ow->ptrs[i], ow->bytes[i])) cudaFree(ow->ptrs[i]);
        ow->ptrs[i] = nullptr;
    }
    *ow_out = ow; *d_out = d_uniq; *count = uniq;
    return BLG_OK;
}
}  // namespace

int blg_sorted_unique_kmers(blg_ctx* c, const char* bases, size_t n, const uint64_t* offs, size_t n_seq, const blg_params* p, blg_kmer_out* out)
{
    if (!c || !p || !out) return BLG_E_INVALID;
    Timer tm;
    blg_stats acc{};
    Owner* ow = nullptr;
    unsigned long long* d = nullptr;
    uint64_t cnt = 0;
    int rc = sorted_unique_device(c, bases, n, offs, n_seq, p, &ow, &d, &cnt, &acc);
    if (rc != BLG_OK) return rc;
    out->count = cnt; out->value_hi = nullptr; out->position = nullptr;
    if (p->flags & BLG_F_DEVICE_OUTPUT) { out->value_lo = reinterpret_cast<uint64_t*>(d); out->owner = ow; }
    else {
        Owner* hw = new (std::nothrow) Owner{0, {}, {}, nullptr};
        if (!hw) { free_owner(ow); return BLG_E_NOMEM; }
        if ((rc = alloc_owned(c, false, cnt, &out->value_lo, hw, 0)) != BLG_OK) { free_owner(ow); free_owner(hw); return rc; }
        if (cnt && (cudaMemcpyAsync(out->value_lo, d, cnt * 8, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess || cudaStreamSynchronize(c->stream) != cudaSuccess)) {
            free_owner(ow); free_owner(hw); return BLG_E_CUDA;
        }
        acc.bytes_d2h += cnt * 8;
        free_owner(ow);
        out->owner = hw;
    }
    acc.n_records = cnt;
    acc.total_ms = tm.ms();
    c->stats = acc;
    return BLG_OK;
}

namespace {
// |A u B| of two sorted distinct device lists = |A| + |B| - |A n B|; the intersection is counted by looking one list up in the other
int union_of_sorted(blg_ctx* c, const unsigned long long* d1, uint64_t c1, const unsigned long long* d2, uint64_t c2, int /*bits*/, uint64_t* uni_out,
                    blg_stats* acc)
{
    *uni_out = c1 + c2;
    if (c1 == 0 || c2 == 0) return BLG_OK;
    int rc = ensure(c, c->small, 64);
    if (rc != BLG_OK) return rc;
    unsigned long long* d_cnt = static_cast<unsigned long long*>(c->small.p) + 4;
    unsigned long long inter = 0;
    cudaEventRecord(c->ev0, c->stream);
    bool ok = intersect_count_u64(d1, c1, d2, c2, d_cnt, c->stream) == cudaSuccess;
    cudaEventRecord(c->ev1, c->stream);
    ok = ok && cudaMemcpyAsync(&inter, d_cnt, 8, cudaMemcpyDeviceToHost, c->stream) == cudaSuccess && cudaStreamSynchronize(c->stream) == cudaSuccess;
    if (!ok) { c->err = "jaccard merge failed"; return BLG_E_CUDA; }
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) == cudaSuccess) acc->kernel_ms += ms;
    acc->kernel_launches += 1;
    *uni_out = c1 + c2 - inter;
    return BLG_OK;
}
}  // namespace

int blg_kmer_jaccard(blg_ctx* c, const char* b1, size_t n1, const uint64_t* o1, size_t s1, const char* b2, size_t n2, const uint64_t* o2, size_t s2,
                     const blg_params* p, blg_jaccard_out* out)
{
    if (!c || !p || !out) return BLG_E_INVALID;
    Timer tm;
    blg_stats acc{};
    *out = blg_jaccard_out{};
    Owner *w1 = nullptr, *w2 = nullptr;
    unsigned long long *d1 = nullptr, *d2 = nullptr;
    uint64_t c1 = 0, c2 = 0;
    int rc = sorted_unique_device(c, b1, n1, o1, s1, p, &w1, &d1, &c1, &acc);
    if (rc != BLG_OK) return rc;
    if ((rc = sorted_unique_device(c, b2, n2, o2, s2, p, &w2, &d2, &c2, &acc)) != BLG_OK) { free_owner(w1); return rc; }
    uint64_t uni = 0;
    if ((rc = union_of_sorted(c, d1, c1, d2, c2, 2 * p->k, &uni, &acc)) != BLG_OK) { free_owner(w1); free_owner(w2); return rc; }
    free_owner(w1); free_owner(w2);
    out->size1 = c1; out->size2 = c2; out->union_size = uni; out->intersection = c1 + c2 - uni;
    acc.total_ms = tm.ms();
    c->stats = acc;
    return BLG_OK;
}

// ---- device-resident building blocks of the multi-GPU form of the sort / unique / Jaccard row (biolib_b200/multi_gpu.py) ----------------
int blg_sort_unique_u64(blg_ctx* c, uint64_t* d_keys, size_t n, int key_bits, uint64_t* count)
{
    if (!c || !count || (n && !d_keys) || key_bits < 1 || key_bits > 64) return BLG_E_INVALID;
    Timer tm;
    c->stats = blg_stats{};
    c->err.clear();
    *count = 0;
    CK(c, cudaSetDevice(c->device));
    if (n == 0) return BLG_OK;
    int rc;
    Owner* ow = new (std::nothrow) Owner{1, {}, {}, nullptr};
    auto fail = [&](int code) { free_owner(ow); return code; };
    if (!ow) return BLG_E_NOMEM;
    unsigned long long *tmp = nullptr, *flags = nullptr, *uq = nullptr;
    if ((rc = alloc_owned(c, true, n, &tmp, ow, 0)) != BLG_OK || (rc = alloc_owned(c, true, n, &uq, ow, 2)) != BLG_OK) return fail(rc);
    if ((rc = ensure(c, c->rank, sort_scratch_bytes(n) + 64)) != BLG_OK) return fail(rc);
    unsigned long long* d_tot = reinterpret_cast<unsigned long long*>(static_cast<char*>(c->rank.p) + sort_scratch_bytes(n) / 8 * 8);
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(d_keys);
    int launches = 0;
    cudaEventRecord(c->ev0, c->stream);
    bool ok = radix_sort_u64(keys, tmp, n, key_bits, c->rank.p, c->stream, &launches) == cudaSuccess;
    ok = ok && unique_sorted_u64(keys, n, flags, c->rank.p, uq, d_tot, c->stream, &launches) == cudaSuccess;
    cudaEventRecord(c->ev1, c->stream);
    unsigned long long u = 0;
    ok = ok && cudaMemcpyAsync(&u, d_tot, 8, cudaMemcpyDeviceToHost, c->stream) == cudaSuccess && cudaStreamSynchronize(c->stream) == cudaSuccess;
    ok = ok && (u == 0 || cudaMemcpyAsync(keys, uq, u * 8, cudaMemcpyDeviceToDevice, c->stream) == cudaSuccess) && cudaStreamSynchronize(c->stream) == cudaSuccess;
    if (!ok) { c->err = "sort / unique failed"; return fail(BLG_E_CUDA); }
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) == cudaSuccess) c->stats.kernel_ms = ms;
    c->stats.kernel_launches = launches;
    c->stats.n_records = u;
    free_owner(ow);
    *count = u;
    c->stats.total_ms = tm.ms();
    return BLG_OK;
}

int blg_sorted_sets_jaccard(blg_ctx* c, const uint64_t* d_a, size_t na, const uint64_t* d_b, size_t nb, int key_bits, blg_jaccard_out* out)
{
    if (!c || !out || (na && !d_a) || (nb && !d_b) || key_bits < 1 || key_bits > 64) return BLG_E_INVALID;
    Timer tm;
    blg_stats acc{};
    c->err.clear();
    *out = blg_jaccard_out{};
    CK(c, cudaSetDevice(c->device));
    uint64_t uni = 0;
    int rc = union_of_sorted(c, reinterpret_cast<const unsigned long long*>(d_a), na, reinterpret_cast<const unsigned long long*>(d_b), nb, key_bits, &uni, &acc);
    if (rc != BLG_OK) return rc;
    out->size1 = na; out->size2 = nb; out->union_size = uni; out->intersection = na + nb - uni;
    acc.total_ms = tm.ms();
    c->stats = acc;
    return BLG_OK;
}

int blg_partition_by_top_byte(blg_ctx* c, const uint64_t* d_keys, size_t n, int key_bits, uint64_t* d_out, uint64_t* bin_counts)
{
    if (!c || !bin_counts || (n && (!d_keys || !d_out)) || key_bits < 1 || key_bits > 64) return BLG_E_INVALID;
    Timer tm;
    c->stats = blg_stats{};
    c->err.clear();
    CK(c, cudaSetDevice(c->device));
    for (int i = 0; i < 256; ++i) bin_counts[i] = 0;
    if (n == 0) return BLG_OK;
    int rc = ensure(c, c->rank, sort_scratch_bytes(n) + 64 + 257 * 8);
    if (rc != BLG_OK) return rc;
    unsigned long long* d_start = reinterpret_cast<unsigned long long*>(static_cast<char*>(c->rank.p) + sort_scratch_bytes(n) / 8 * 8 + 64);
    unsigned long long start[257];
    int launches = 0;
    CK(c, cudaEventRecord(c->ev0, c->stream));
    CK(c, partition_by_digit_u64(reinterpret_cast<const unsigned long long*>(d_keys), reinterpret_cast<unsigned long long*>(d_out), n,
                                 key_bits > 8 ? key_bits - 8 : 0, c->rank.p, d_start, c->stream, &launches));
    CK(c, cudaEventRecord(c->ev1, c->stream));
    CK(c, cudaMemcpyAsync(start, d_start, sizeof(start), cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    for (int i = 0; i < 256; ++i) bin_counts[i] = start[i + 1] - start[i];
    add_kernel_time(c);
    c->stats.kernel_launches = launches;
    c->stats.n_records = n;
    c->stats.total_ms = tm.ms();
    return BLG_OK;
}

int blg_lower_bound_u64(blg_ctx* c, const uint64_t* d_sorted, size_t n, const uint64_t* queries, size_t nq, uint64_t* out_index)
{
    if (!c || (n && !d_sorted) || (nq && (!queries || !out_index))) return BLG_E_INVALID;
    c->err.clear();
    CK(c, cudaSetDevice(c->device));
    if (nq == 0) return BLG_OK;
    int rc = ensure(c, c->rank, 2 * nq * 8);
    if (rc != BLG_OK) return rc;
    unsigned long long* d_q = static_cast<unsigned long long*>(c->rank.p);
    CK(c, cudaMemcpyAsync(d_q, queries, nq * 8, cudaMemcpyHostToDevice, c->stream));
    CK(c, lower_bound_u64(reinterpret_cast<const unsigned long long*>(d_sorted), n, d_q, nq, d_q + nq, c->stream));
    CK(c, cudaMemcpyAsync(out_index, d_q + nq, nq * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    return BLG_OK;
}

int blg_hash64(blg_ctx* c, const uint64_t* values, size_t n, uint64_t seed, uint32_t flags, uint64_t* out)
{
    if (!c || (n && (!values || !out))) return BLG_E_INVALID;
    c->err.clear();
    CK(c, cudaSetDevice(c->device));
    const unsigned long long s8 = (unsigned long long)(uint32_t)seed ^ 8ULL;
    if (flags & (BLG_F_DEVICE_INPUT | BLG_F_DEVICE_OUTPUT)) {
        CK(c, hash64_launch(reinterpret_cast<const unsigned long long*>(values), reinterpret_cast<unsigned long long*>(out), n, s8, c->stream));
        CK(c, cudaStreamSynchronize(c->stream));
        return BLG_OK;
    }
    int rc;
    if ((rc = ensure(c, c->o_value, n * 8)) != BLG_OK || (rc = ensure(c, c->o_mm, n * 8)) != BLG_OK) return rc;
    if (n) {
        CK(c, cudaMemcpyAsync(c->o_value.p, values, n * 8, cudaMemcpyHostToDevice, c->stream));
        CK(c, hash64_launch(static_cast<unsigned long long*>(c->o_value.p), static_cast<unsigned long long*>(c->o_mm.p), n, s8, c->stream));
        CK(c, cudaMemcpyAsync(out, c->o_mm.p, n * 8, cudaMemcpyDeviceToHost, c->stream));
        CK(c, cudaStreamSynchronize(c->stream));
    }
    return BLG_OK;
}

int blg_minimizer_positions(blg_ctx* c, const uint64_t* values, size_t n, uint8_t k, uint8_t m, uint32_t flags, uint8_t* out)
{
    if (!c || (n && (!values || !out))) return BLG_E_INVALID;
    c->err.clear();
    if (k == 0 || k > 32 || m == 0 || m > k) { c->err = "need 1 <= m <= k <= 32 (k-mer values held in uint64_t)"; return BLG_E_INVALID; }
    CK(c, cudaSetDevice(c->device));
    if (flags & (BLG_F_DEVICE_INPUT | BLG_F_DEVICE_OUTPUT)) {
        CK(c, minpos_launch(reinterpret_cast<const unsigned long long*>(values), n, k, m, out, c->sm_count, c->stream));
        CK(c, cudaStreamSynchronize(c->stream));
        return BLG_OK;
    }
    int rc;
    if ((rc = ensure(c, c->o_value, n * 8)) != BLG_OK || (rc = ensure(c, c->o_size, n)) != BLG_OK) return rc;
    if (n) {
        CK(c, cudaMemcpyAsync(c->o_value.p, values, n * 8, cudaMemcpyHostToDevice, c->stream));
        CK(c, minpos_launch(static_cast<unsigned long long*>(c->o_value.p), n, k, m, static_cast<unsigned char*>(c->o_size.p), c->sm_count, c->stream));
        CK(c, cudaMemcpyAsync(out, c->o_size.p, n, cudaMemcpyDeviceToHost, c->stream));
        CK(c, cudaStreamSynchronize(c->stream));
    }
    return BLG_OK;
}

void blg_free_kmer_out(blg_ctx* c, blg_kmer_out* out)
{
    if (!out) return;
    if (c) cudaSetDevice(c->device);
    if (out->owner) {
        free_owner(static_cast<Owner*>(out->owner));
        out->value_lo = out->value_hi = out->position = nullptr;
        out->run_first_record = out->run_start = nullptr;
        out->n_runs = 0;
        out->owner = nullptr;
        out->count = 0;
    }
}

void blg_free_sk_out(blg_ctx* c, blg_sk_out* out)
{
    if (!out) return;
    if (c) cudaSetDevice(c->device);
    if (out->owner) {
        free_owner(static_cast<Owner*>(out->owner));
        out->value = out->mm_start = out->hash = out->id = nullptr;
        out->size = out->mm_pos = nullptr;
        out->run_first_record = out->run_start = nullptr;
        out->n_runs = 0;
        out->value_lo32 = nullptr; out->value_hi8 = out->size_pos = nullptr;
        out->owner = nullptr;
        out->count = 0;
    }
}

}  // extern "C"
