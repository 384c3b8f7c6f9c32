/*
 * biolib_gpu.h -- C ABI of the B200 (sm_100a) streaming k-mer path.
 *
 * This is the drop-in boundary for ONE path of yhhshb/biolib: ASCII DNA -> 2-bit codes ->
 * k-mers / m-mer hashes -> window minimizers -> super-k-mers. The reference has no FFI or
 * plugin layer for it; its boundary is three header-only class templates, and each entry point
 * below replaces one of them in batched form (file:line relative to the reference tree):
 *
 *   blg_kmers        <- wrapper::kmer_view<KmerType,Iterator>(start, stop, k, canonical)
 *                       include/kmer_view.hpp:61, iteration :155-234
 *   blg_minimizers   <- wrapper::minimizer_view<K,M,Hash,Iterator>(start, stop, k, m, seed, canonical)
 *                       include/minimizer_view.hpp:88, iteration :166-266, hash include/hash.hpp:55-59
 *   blg_super_kmers  <- wrapper::super_kmer_view<K,M,Hash>(contig, len, k, m, canonical)
 *                       include/super_kmer_view.hpp:48-49, iteration :104-135
 *   blg_hash64       <- hash::hash64::hash<uint64_t>(val, seed)          include/hash.hpp:55-59
 *   blg_minimizer_positions <- hash::minimizer_position_extractor::operator()   include/kmer_view.hpp:266-283
 *
 * The C++ shim that keeps the reference's constructor surface over these calls is
 * include/biolib_gpu.hpp. Plain pointers and sizes only; no exceptions cross this boundary;
 * every function returns a blg_status. There is no CPU fallback: without a CUDA device every
 * compute call fails with BLG_E_CUDA.
 *
 * Coordinates. `bases` is one buffer of n_bases ASCII bytes. `seq_offsets` (n_seq + 1 entries,
 * offsets[0] == 0, offsets[n_seq] == n_bases, non-decreasing) splits it into independent
 * sequences; NULL means one sequence. All positions returned are 0-based offsets into `bases`
 * (the shim rebases them per sequence and applies the reference's +1).
 * Nothing spans two sequences, and nothing spans a byte that is not one of AaCcGgTtUu
 * (include/constants.hpp:12-21).
 */
#ifndef BIOLIB_GPU_H
#define BIOLIB_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BLG_ABI_VERSION 3

typedef enum blg_status {
    BLG_OK = 0,
    BLG_E_INVALID = 1,     /* bad argument: k == 0, k > 64, m == 0, m > 32, m > k, bad offsets ... */
    BLG_E_UNSUPPORTED = 2, /* a combination this build does not implement */
    BLG_E_CAPACITY = 3,    /* caller-provided output too small; out->count holds the required size */
    BLG_E_CUDA = 4,        /* CUDA runtime error or no device; see blg_last_error() */
    BLG_E_NOMEM = 5
} blg_status;

/* flags */
#define BLG_F_DEVICE_INPUT  0x1u /* bases / seq_offsets are device pointers on the context's GPU */
#define BLG_F_DEVICE_OUTPUT 0x2u /* output arrays live in device memory (caller- or library-owned) */
#define BLG_F_EMIT_HASH     0x4u /* also fill out->hash (parity / debugging; not algorithmic bytes) */
#define BLG_F_NO_SIZE       0x8u /* blg_minimizers sets it: skip size / mm_pos (16 B records) */
#define BLG_F_POS32         0x20u /* compact records: out->mm_start points at uint32_t entries holding mm_start - win_begin (range
                                   * form; mm_start itself for the whole-buffer calls). 14 instead of 18 bytes per record over PCIe.
                                   * BLG_E_INVALID when a position could exceed 2^32 - 1; not combinable with BLG_F_EMIT_ID */
#define BLG_F_REF_DROP_LAST 0x40u /* blg_sorted_unique_kmers / blg_kmer_jaccard: leave out the last k-mer of every sequence that ends in
                                   * a valid base, as iterating the reference's kmer_view does (include/kmer_view.hpp:57) */
#define BLG_F_ASYNC         0x80u /* with BLG_F_DEVICE_INPUT | BLG_F_DEVICE_OUTPUT and caller arrays: only enqueue the kernel on the context's
                                   * stream and return; out->count stays 0. The record count (and a BLG_E_CAPACITY verdict) comes from
                                   * blg_async_count(), which waits for the stream. Several calls may be in flight; the count is the last one's */
#define BLG_F_RUNS          0x100u /* (blg_kmers with host output: see blg_kmer_out.) Smallest records over PCIe (host output of the super-k-mer calls): mm_start is NOT returned (out->mm_start stays
                                    * NULL / untouched); instead the records come with the list of RUNS -- maximal stretches of records in which
                                    * every record starts where the previous one ended (one run for a sequence without breaks):
                                    *   run_first_record[j]  index of the first record of run j      run_start[j]  its sk_start
                                    * so that sk_start[r] = run_start[j] + sum of size[run_first_record[j] .. r), mm_start = sk_start + mm_pos.
                                    * 10 instead of 18 bytes per record. With caller arrays set run_capacity and the two run arrays. */
#define BLG_F_TIGHT         0x200u /* with BLG_F_RUNS, for m <= 20 and k - m <= 15: the same records in 6 bytes (5 for m <= 16) instead of 10 over PCIe.
                                    * out->value / size / mm_pos stay NULL; instead
                                    *   value_lo32[i]  bits 0..31 of the minimizer value      value_hi8[i]  bits 32..39 (NULL / untouched for m <= 16)
                                    *   size_pos[i]    (size - 1) | mm_pos << 4
                                    * (packed on the device by one more small kernel per piece). Not combinable with BLG_F_EMIT_HASH / _ID. */
#define BLG_F_EMIT_ID       0x10u /* also fill out->id (minimizer_context_t::id; one extra pass over the bases) */

typedef struct blg_params {
    uint8_t k;          /* k-mer length, 1..64 (kmers), m..64 (minimizers / super-k-mers) */
    uint8_t m;          /* minimizer length, 1..32 (MinimizerType = uint64_t); unused by blg_kmers */
    uint8_t canonical;  /* 0 / 1: value = min(fwd, reverse complement) */
    uint8_t reserved;
    uint32_t flags;     /* BLG_F_* */
    uint64_t seed;      /* narrowed to 32 bits exactly as include/hash.hpp:50 does */
} blg_params;

/*
 * Output of blg_kmers: one entry per valid k-mer, in position order.
 *   value_lo[i] (and value_hi[i] when k > 32; may be NULL for k <= 32), position[i].
 * The reference's exact record stream (nullopt records after breaks, ids, the dropped last
 * k-mer: SURVEY K1-K4) is rebuilt from `position` by the shim.
 */
typedef struct blg_kmer_out {
    uint64_t count;     /* out: number of records (required size on BLG_E_CAPACITY) */
    uint64_t capacity;  /* in: entries available in caller arrays; 0 = library allocates */
    uint64_t* value_lo;
    uint64_t* value_hi;
    uint64_t* position;
    void* owner;        /* library bookkeeping for blg_free_*; leave NULL */
    /* only with BLG_F_RUNS (blg_kmers with host output; ABI version 3): `position` is NOT returned (stays NULL / untouched). The
     * k-mers of a stretch of valid bases start at consecutive positions, so the list of those stretches says where every k-mer is:
     *   position[i] = run_start[j] + (i - run_first_record[j])   for run_first_record[j] <= i < run_first_record[j + 1]
     * 8 instead of 16 bytes per k-mer over PCIe (16 instead of 24 for k > 32): the dense layout of SURVEY 8(d). */
    uint64_t n_runs;            /* out (required size on BLG_E_CAPACITY) */
    uint64_t run_capacity;      /* in, with caller arrays */
    uint64_t* run_first_record;
    uint64_t* run_start;
} blg_kmer_out;

/*
 * Output of blg_minimizers / blg_super_kmers: one record per maximal run of consecutive valid
 * k-mers sharing the same minimizer occurrence, in order of first k-mer.
 *   value[i]    canonical-or-forward 2-bit m-mer (minimizer_context_t::value, minimizer_view.hpp:22)
 *   mm_start[i] 0-based start of the minimizer in `bases`
 *   size[i]     number of k-mers in the run, 1..k-m+1       (super_kmer_t::size, super_kmer_view.hpp:23)
 *   mm_pos[i]   mm_start - (start of the first k-mer), 0..k-m (super_kmer_t::mm_pos, :22)
 *   hash[i]     hash64(value, seed), only with BLG_F_EMIT_HASH
 *   id[i]       number of valid m-mer starts before mm_start inside the same sequence (minimizer_view.hpp:257,263),
 *               only with BLG_F_EMIT_ID
 * blg_minimizers leaves size / mm_pos untouched (may be NULL).
 */
typedef struct blg_sk_out {
    uint64_t count;
    uint64_t capacity;
    uint64_t* value;
    uint64_t* mm_start;
    uint8_t* size;
    uint8_t* mm_pos;
    uint64_t* hash;
    uint64_t* id;       /* only with BLG_F_EMIT_ID: ordinal of the minimizer's m-mer among the valid m-mers of its sequence */
    void* owner;
    /* only with BLG_F_RUNS */
    uint64_t n_runs;            /* out (required size on BLG_E_CAPACITY) */
    uint64_t run_capacity;      /* in, with caller arrays */
    uint64_t* run_first_record;
    uint64_t* run_start;
    /* only with BLG_F_TIGHT (ABI version 3) */
    uint32_t* value_lo32;
    uint8_t* value_hi8;
    uint8_t* size_pos;
} blg_sk_out;

typedef struct blg_stats {
    uint64_t n_bases;        /* bases of the last call */
    uint64_t n_records;      /* records produced */
    uint64_t bytes_h2d;      /* bytes copied host -> device by the last call */
    uint64_t bytes_d2h;      /* bytes copied device -> host by the last call */
    uint64_t kernel_launches;/* kernels of this library launched by the last call */
    double kernel_ms;        /* device time of those kernels (CUDA events on the context stream) */
    double total_ms;         /* wall time of the call */
    uint64_t exact_tiles;    /* tiles of the fused kernel redone with full 64-bit hash compares (ties in the 32-bit keys) */
} blg_stats;

typedef struct blg_ctx blg_ctx;

/* context: one GPU, one stream; not thread-safe (one context per host thread / GPU) */
int blg_create(int device, blg_ctx** out);
int blg_destroy(blg_ctx* ctx);
/* Run on a caller stream (cudaStream_t passed as void*); NULL (which is also the value of the legacy default stream 0: use a
 * created stream) restores the context's own non-blocking stream.
 * Device-resident calls (BLG_F_DEVICE_INPUT|BLG_F_DEVICE_OUTPUT with caller arrays) then only
 * enqueue work and one 8-byte count read-back. */
int blg_set_stream(blg_ctx* ctx, void* cuda_stream);
/* Shape of the chunked host path (inputs of 8 Mi k-mer starts and more with host buffers: upload, kernel and download of successive
 * pieces overlap): pieces of chunk_mi Mi k-mer starts (at most 16 pieces per call), the first piece 1 / first_div of a regular one so
 * that the download starts early. The defaults are tuned on a PCIe Gen5 x16 B200 box; results never depend on them. */
int blg_set_pipeline(blg_ctx* ctx, int chunk_mi, int first_div);
int blg_abi_version(void);
const char* blg_strerror(int status);
const char* blg_last_error(blg_ctx* ctx);
int blg_last_stats(blg_ctx* ctx, blg_stats* out);
/* waits for the context's stream; *count = records of the last BLG_F_ASYNC call (BLG_E_CAPACITY if they exceeded its capacity) */
int blg_async_count(blg_ctx* ctx, uint64_t* count);
/* number of SMs / device name of the context's GPU (for sizing and reports) */
int blg_device_info(blg_ctx* ctx, int* sm_count, char* name, size_t name_len);

/* pinned host memory for callers that want full-speed H2D / D2H */
int blg_host_alloc(blg_ctx* ctx, size_t bytes, void** out);
int blg_host_free(blg_ctx* ctx, void* p);

int blg_kmers(blg_ctx* ctx, const char* bases, size_t n_bases, const uint64_t* seq_offsets, size_t n_seq,
              const blg_params* p, blg_kmer_out* out);
int blg_minimizers(blg_ctx* ctx, const char* bases, size_t n_bases, const uint64_t* seq_offsets, size_t n_seq,
                   const blg_params* p, blg_sk_out* out);
int blg_super_kmers(blg_ctx* ctx, const char* bases, size_t n_bases, const uint64_t* seq_offsets, size_t n_seq,
                    const blg_params* p, blg_sk_out* out);

/*
 * Range form used by the multi-GPU driver: emit only the records whose FIRST k-mer starts in
 * [win_begin, win_end) (positions in `bases` coordinates); the call reads the k+w-2 halo it needs
 * from `bases` itself, so ranks given the same buffer (or a slice with halos, rebased by the
 * caller) produce disjoint, concatenable outputs.
 */
int blg_super_kmers_range(blg_ctx* ctx, const char* bases, size_t n_bases, const uint64_t* seq_offsets, size_t n_seq,
                          uint64_t win_begin, uint64_t win_end, const blg_params* p, blg_sk_out* out);

/* batched hash::hash64::hash<uint64_t> (host pointers unless BLG_F_DEVICE_INPUT|OUTPUT semantics
 * are requested through `flags`: both arrays then are device pointers) */
int blg_hash64(blg_ctx* ctx, const uint64_t* values, size_t n, uint64_t seed, uint32_t flags, uint64_t* out);

/*
 * batched hash::minimizer_position_extractor::operator() (include/kmer_view.hpp:266-283; the PropertyExtractor that
 * sampler::syncmer_sampler is built on): for every k-mer value (2-bit packed, k <= 32, as blg_kmers returns them)
 * the 0-based offset, from the left end of the k-mer, of its minimum-hash m-mer (hash64 with seed 0 of the plain
 * m-mer bits; leftmost on ties). Host pointers unless `flags` has BLG_F_DEVICE_INPUT|BLG_F_DEVICE_OUTPUT (both arrays
 * then are device pointers). The reference's nullopt branch (returns k + 1) belongs to the shim: values here are plain.
 */
int blg_minimizer_positions(blg_ctx* ctx, const uint64_t* kmer_values, size_t n, uint8_t k, uint8_t m, uint32_t flags,
                            uint8_t* out_positions);

/*
 * sampler::syncmer_sampler<kmer_view::const_iterator, hash::minimizer_position_extractor>
 * (include/syncmer_sampler.hpp:130-137 over include/kmer_view.hpp:266-283), fused into the k-mer kernel: the entries
 * blg_kmers would return, restricted to the k-mers whose minimizer_position_extractor(k, p->m) offset equals
 * start_offset or end_offset (closed syncmers: 0 and k - m). The offset is that of the minimum hash64(seed 0) among the
 * plain m-mer bit fields of the (canonical, if p->canonical) k-mer VALUE, leftmost on ties - not of the minimizer_view
 * m-mers. 1 <= p->m <= p->k <= 32; p->seed is unused. The reference's record stream drops the last k-mer of a sequence
 * that ends in a valid base and its nullopt records never pass the filter: the shim applies K1 as for blg_kmers.
 */
int blg_syncmers(blg_ctx* ctx, const char* bases, size_t n_bases, const uint64_t* seq_offsets, size_t n_seq,
                 const blg_params* p, uint16_t start_offset, uint16_t end_offset, blg_kmer_out* out);

/*
 * sampler::hash_sampler<Iterator, hash::hash64> over the k-mer values of a kmer_view (include/hash_sampler.hpp:73-78,
 * 138-141), fused into the k-mer kernel: the entries blg_kmers would return, restricted to those with
 * hash64(value, p->seed) < threshold. `threshold` is the reference's
 * static_cast<uint64_t>(sampling_rate * std::numeric_limits<uint64_t>::max()) (hash_sampler.hpp:77), computed by the
 * caller so that the double arithmetic stays on the reference's side. k <= 32 (64-bit values); p->m is unused.
 */
int blg_hash_sampled_kmers(blg_ctx* ctx, const char* bases, size_t n_bases, const uint64_t* seq_offsets, size_t n_seq,
                           const blg_params* p, uint64_t threshold, blg_kmer_out* out);

/*
 * blg_kmers for callers that may not allocate. No counterpart in the reference: added for code
 * that cannot call cudaMalloc (under CUDA graph capture, in tight serving loops, in harnesses that forbid it). The same kernel and
 * the same (value, position) records as blg_kmers with BLG_F_DEVICE_INPUT | BLG_F_DEVICE_OUTPUT, one sequence, but context-free: it
 * runs on the CURRENT device, on `stream` (a cudaStream_t; NULL = the legacy default stream), and its scan scratch is a static
 * device array of the library (129 KB per GPU that loads the library, held until exit) instead of cudaMalloc'ed buffers. It calls
 * no cudaMalloc, cudaFree or allocating library routine.
 *   limit     n_bases <= 67108864 (2^26); beyond it BLG_E_CAPACITY (and the text in err). Longer inputs: blg_kmers, or several calls
 *   inputs    d_bases (device, ASCII), k 1..64, canonical 0 / 1
 *   outputs   d_value_lo[capacity], d_value_hi[capacity] (k > 32 only, else may be NULL), d_position[capacity] (may be NULL): device
 *             memory of the caller; at most n_bases - k + 1 records
 *   count     d_count (device, 8 bytes) and / or count (host). With `count` the call waits for the stream and returns BLG_E_CAPACITY if
 *             the records did not fit (*count = required size); with d_count alone it only enqueues. One of the two is required
 *   err       optional buffer for the error text (err_len bytes)
 * All calls on a device share the one scratch array; the call clears what must start at zero on `stream` first. Calls that overlap on
 * the GPU (different streams, or different threads without ordering between their streams) race on it: serialise them.
 */
int blg_kmers_static_scratch(const char* d_bases, size_t n_bases, uint8_t k, int canonical, uint64_t* d_value_lo, uint64_t* d_value_hi,
                             uint64_t* d_position, uint64_t capacity, void* stream, uint64_t* d_count, uint64_t* count, char* err,
                             size_t err_len);

/*
 * The reference's one downstream pipeline (tests/test_jaccard.cpp:55-130) on the device, k <= 32:
 *   blg_sorted_unique_kmers  kmer_view values -> emem::external_memory_vector (sorted, include/external_memory_vector.hpp:186-196,
 *                            320-341) -> sampler::ordered_unique_sampler (include/ordered_unique_sampler.hpp:116-130): the distinct
 *                            k-mer values of the input in increasing order, in out->value_lo (position / value_hi stay NULL)
 *   blg_kmer_jaccard         that for two inputs + algorithm::jaccard (include/jaccard.hpp:8-38): {intersection, union, size1, size2}
 * p->k, p->canonical, p->flags (BLG_F_DEVICE_INPUT, BLG_F_DEVICE_OUTPUT for the sorted list, BLG_F_REF_DROP_LAST) apply.
 */
typedef struct blg_jaccard_out {
    uint64_t intersection;
    uint64_t union_size;
    uint64_t size1;
    uint64_t size2;
} blg_jaccard_out;
int blg_sorted_unique_kmers(blg_ctx* ctx, const char* bases, size_t n_bases, const uint64_t* seq_offsets, size_t n_seq,
                            const blg_params* p, blg_kmer_out* out);
int blg_kmer_jaccard(blg_ctx* ctx, const char* bases1, size_t n1, const uint64_t* seq_offsets1, size_t n_seq1,
                     const char* bases2, size_t n2, const uint64_t* seq_offsets2, size_t n_seq2, const blg_params* p,
                     blg_jaccard_out* out);

/*
 * The same row across several GPUs (SURVEY.md 8(e) + 8(f) rank 3; biolib_b200/multi_gpu.py drives it, one process per GPU): every
 * rank extracts, sorts and deduplicates its share of the input (blg_sorted_unique_kmers with BLG_F_DEVICE_OUTPUT), the ranks agree on
 * key-range splitters, exchange the ranges (one all-to-all over NVLink: the path's one real exchange step), and each rank merges what it
 * received. These are the device-resident pieces of that, all on the context's stream, all pointers device pointers unless stated:
 *   blg_sort_unique_u64      d_keys[0, n) -> sorted distinct keys in d_keys[0, *count) (in place). key_bits = number of low bits the
 *                            keys can differ in (2k for k-mer values)
 *   blg_sorted_sets_jaccard  algorithm::jaccard (include/jaccard.hpp:8-38) of two sorted distinct lists
 *   blg_partition_by_top_byte  d_keys[0, n) -> d_out grouped (stably) by their top byte, (key >> (key_bits - 8)) & 255, groups in increasing
 *                            order; bin_counts (HOST, 256 entries) = keys per group. The partition step when the keys are exchanged
 *                            BEFORE they are sorted: a rank's share of the key space is a range of groups, i.e. a contiguous piece of d_out
 *   blg_lower_bound_u64      for each of nq HOST queries the index of the first key >= it in the sorted device array (HOST out_index):
 *                            where a rank's sorted list is cut for the exchange
 */
int blg_sort_unique_u64(blg_ctx* ctx, uint64_t* d_keys, size_t n, int key_bits, uint64_t* count);
int blg_sorted_sets_jaccard(blg_ctx* ctx, const uint64_t* d_a, size_t na, const uint64_t* d_b, size_t nb, int key_bits, blg_jaccard_out* out);
int blg_partition_by_top_byte(blg_ctx* ctx, const uint64_t* d_keys, size_t n, int key_bits, uint64_t* d_out, uint64_t* bin_counts);
int blg_lower_bound_u64(blg_ctx* ctx, const uint64_t* d_sorted, size_t n, const uint64_t* queries, size_t nq, uint64_t* out_index);

/*
 * Super-k-mer payloads grouped by minimizer bucket (what super-k-mers are for -- partitioned k-mer
 * counting, de Bruijn graph construction). The records of blg_super_kmers, reordered STABLY by
 *   bucket = hash64(value, p->seed) % n_buckets          (hash::hash64::hash, include/hash.hpp:55-59)
 * together with the bases [sk_start, sk_start + size + k - 1) of every record appended, in that order, to one 2-bit packed
 * stream in the layout of the reference's bit::packed::vector<uint64_t>(2) (include/packed_vector.hpp:134-160,311-318:
 * element i in words[i / 32] at bits [62 - 2 (i % 32), 64 - 2 (i % 32)), A=0 C=1 G=2 T/U=3): n_words = total_bases * 2 / 64 + 1.
 *   bucket_offsets[b] .. bucket_offsets[b+1]   records of bucket b   (n_buckets + 1 entries)
 *   payload_start[i]                            index of record i's first base in the stream; it holds size[i] + k - 1 bases
 * 1 <= n_buckets <= 4096. flags: BLG_F_DEVICE_INPUT, BLG_F_DEVICE_OUTPUT. Arrays are library-allocated (pinned host or device);
 * release with blg_free_payload_out.
 */
typedef struct blg_payload_out {
    uint64_t count;
    uint32_t n_buckets;
    uint64_t total_bases;
    size_t n_words;
    uint64_t* bucket_offsets;
    uint64_t* value;
    uint64_t* mm_start;
    uint8_t* size;
    uint8_t* mm_pos;
    uint64_t* payload_start;
    uint64_t* words;
    void* owner;
} blg_payload_out;
int blg_super_kmer_payloads(blg_ctx* ctx, const char* bases, size_t n_bases, const uint64_t* seq_offsets, size_t n_seq,
                            const blg_params* p, uint32_t n_buckets, blg_payload_out* out);
void blg_free_payload_out(blg_ctx* ctx, blg_payload_out* out);

/*
 * Ingest: the two input formats either side of the hot path.
 *
 * 2-bit packed input. H2D then moves 0.25 B per base. `codes` holds 16 bases per 32-bit little-endian word, base q of a
 * word at bits [2q, 2q+2) with A=0 C=1 G=2 T/U=3 (the order of constants::seq_nt4_table, include/constants.hpp:12-21);
 * ceil(n_bases / 16) words. Everything the table maps to 4 (N, IUPAC codes, ...) is listed as break intervals
 * [break_start[i], break_start[i] + break_len[i]), sorted, disjoint; the codes under a break are ignored. Case is not kept:
 * no output depends on it.
 *   blg_pack_bases          host-side packer: ASCII -> codes + break intervals (BLG_E_CAPACITY with *n_breaks = the number
 *                           needed when break_capacity is too small)
 *   blg_super_kmers_packed  blg_super_kmers_range over packed input (host pointers); the bases are unpacked on the device
 */
typedef struct blg_packed_seq {
    const uint32_t* codes;
    size_t n_bases;
    const uint64_t* break_start;
    const uint32_t* break_len;
    size_t n_breaks;
} blg_packed_seq;
int blg_pack_bases(const char* bases, size_t n_bases, uint32_t* codes, uint64_t* break_start, uint32_t* break_len,
                   size_t break_capacity, size_t* n_breaks);
int blg_super_kmers_packed(blg_ctx* ctx, const blg_packed_seq* seq, const uint64_t* seq_offsets, size_t n_seq,
                           uint64_t win_begin, uint64_t win_end, const blg_params* p, blg_sk_out* out);

/*
 * FASTA / FASTQ text -> concatenated sequence bytes + seq_offsets (n_seq + 1 entries), i.e. the (bases, seq_offsets) pair
 * every entry point above takes: what the reference obtains record by record from kseq_read() (tests/kseq.h:185-240, as used
 * by tests/test_kmer_view.cpp:30-40), split on the device. format: 0 = FASTA (multi-line records), 1 = FASTQ (four lines per
 * record), -1 = by the first byte ('>' / '@'). '\n' and '\r' are dropped; every other byte of a sequence line is kept.
 * flags: BLG_F_DEVICE_INPUT (text is a device pointer, 16-byte aligned), BLG_F_DEVICE_OUTPUT (bases / seq_offsets stay on the
 * device, library-owned: feed them to the entry points above with BLG_F_DEVICE_INPUT). Release with blg_free_fastx_out.
 * BLG_E_INVALID when the text does not start with a record header.
 */
typedef struct blg_fastx_out {
    uint64_t n_bases;
    uint64_t n_seq;
    char* bases;
    uint64_t* seq_offsets;
    void* owner;
} blg_fastx_out;
int blg_parse_fastx(blg_ctx* ctx, const char* text, size_t n_text, int format, uint32_t flags, blg_fastx_out* out);
void blg_free_fastx_out(blg_ctx* ctx, blg_fastx_out* out);

/*
 * Host-side workload generators (csrc/synth_host.cpp). They run on the CPU, as the
 * reference's own do, and feed the benchmarks / tests; nothing on the GPU path calls them.
 *   blg_synth_generate <- sequence::generator::nucleic<packed::uniform>::get_sequence(length) with uniform(seed)
 *                         include/sequence_generator.hpp:14-23,28-43
 *   blg_synth_mutate   <- wrapper::sequence_mutator(src, src + n, mutp, indelf, extp).cbegin(seed) iterated to cend()
 *                         with the absent values dropped (include/sequence_mutator.hpp:109-169, consumed as in
 *                         tests/test_random_sequence_generation.cpp:44-48). *out_len = bases produced (the required
 *                         size on BLG_E_CAPACITY); counts = {substitutions, insertions, deletions}, may be NULL.
 */
int blg_synth_generate(uint64_t seed, size_t length, char* out);
int blg_synth_mutate(const char* src, size_t n, double mutp, double indelf, double extp, uint64_t seed, char* out,
                     size_t capacity, size_t* out_len, uint64_t counts[3], int64_t* total_indel_len);

/* release library-allocated output arrays (host pinned or device); caller-provided arrays are untouched */
void blg_free_kmer_out(blg_ctx* ctx, blg_kmer_out* out);
void blg_free_sk_out(blg_ctx* ctx, blg_sk_out* out);

#ifdef __cplusplus
}
#endif
#endif /* BIOLIB_GPU_H */
