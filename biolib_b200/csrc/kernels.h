// kernels.h -- launch interface between the C-ABI layer (api.cu) and the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace blg {

// One launch of the fused minimizer / super-k-mer kernel over windows [win_lo, win_hi).
struct SkArgs {
    const unsigned char* bases;  // address of global position buf_lo
    long long buf_lo, buf_hi;    // positions present in `bases`
    const unsigned long long* seq_offsets;  // device, n_seq + 1 entries, global positions; may be null
    long long n_seq;
    long long win_lo, win_hi;    // emit the records whose first k-mer starts in [win_lo, win_hi)
    int k, m, w;
    int canonical;
    unsigned long long s8;       // (uint64)(uint32)seed ^ 8
    int vec_ok;                  // (bases - buf_lo) is 16-byte aligned
    int rb;                      // real w-blocks per tile = blockDim.x - 3
    int nwords;                  // staged 16-base words per tile (without the 2 pad words)
    long long num_tiles;
    // second-generation kernel (sk2_kernel.cu); filled by sk2_plan
    long long org;               // win_lo rounded down to a multiple of 16: tile grid origin
    int R;                       // real windows per tile (multiple of 16)
    int levels, df;              // doubling levels L = floor(log2 w), last combine offset w - 2^L
    int pp_off[8], pp_entries;   // shared-memory offsets (in 16-byte entries) of the partial level arrays
    int warp_smem;               // bytes of shared memory per warp (each warp owns a tile)
    unsigned long long* value;   // outputs (device); size / mm_pos / hash may be null
    unsigned long long* mm_start;
    unsigned int* mm_start32;    // BLG_F_POS32: positions as uint32 relative to pos_base instead (mm_start is null then)
    long long pos_base;
    unsigned char* size;
    unsigned char* mm_pos;
    unsigned long long* hash;
    unsigned long long capacity;
    unsigned int* tile_counter;        // zeroed before launch
    unsigned long long* tile_status;   // num_tiles entries, zeroed before launch
    unsigned long long* total;         // out: number of records
    // third-generation kernel (sk3_kernel.cu): two-level scan; filled by sk3_plan / the C-ABI layer
    long long status_words;            // u64 words of scan scratch to zero: num_tiles + 2 * groups
    int gq, gr, gl;                    // general form: k - m = gq * B + gr; lanes 1 .. gl own record starts
    unsigned long long* grp_acc;       // [groups] arrivals << 40 | records of a group of 32 tiles
    unsigned long long* grp_status;    // [groups] look-back words over groups
    // fourth-generation kernel (sk4_kernel.cu): 32-bit keys (hash bits above an index field) + exact redo on ties
    unsigned int keymask;              // ~(2^index_bits - 1)
    unsigned long long* exact_tiles;   // optional counter: tiles that took the exact path (may be null); only grows
    unsigned int epoch;                // tag of this launch's scan words (22 bits, never 0): no scratch is zeroed between launches
    unsigned int ticket_base;          // value of *tile_counter when this launch starts (it only grows)
    // syncmer filter in the sk4 frame (sy4_kernel.cu): value = kept k-mer values, mm_start = their positions (may be null)
    int sync_on, sync_s, sync_e;
};

// position store of the record-emitting loops: 64-bit absolute, or 32-bit relative to pos_base (BLG_F_POS32)
#ifdef __CUDACC__
__device__ __forceinline__ void store_mm_start(const SkArgs& a, unsigned long long g, long long pos)
{
    if (a.mm_start32 != nullptr) a.mm_start32[g] = (unsigned int)(pos - a.pos_base);
    else a.mm_start[g] = (unsigned long long)pos;
}
#endif

struct SkLaunchPlan {
    int threads;       // w-blocks per tile (= CTA size)
    size_t smem_bytes;
    int grid;
};

// sk2_kernel.cu: serves k - m < 5 (sparse-table window minimum); fills the launch plan for (k, m) on `sm_count` SMs
#ifndef SK2_THREADS
#define SK2_THREADS 128
#endif
cudaError_t sk2_plan(SkArgs& args, int sm_count, SkLaunchPlan* plan);
cudaError_t sk2_launch(const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream);

// third generation (sk3_kernel.cu): lane-owned blocks of (w-1)/Q starts, van Herk / Gil-Werman minima in registers.
// Covers the (k, m) families of the benchmark configurations; sk3_supported() says which, everything else runs sk2.
#ifndef SK3_THREADS
#define SK3_THREADS 128
#endif
#ifndef SK3_MIN_BLOCKS
#define SK3_MIN_BLOCKS 4
#endif
bool sk3_supported(int k, int m);
cudaError_t sk3_plan(SkArgs& args, int sm_count, SkLaunchPlan* plan);
cudaError_t sk3_launch(const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream);

// fourth generation (sk4_kernel.cu): a lane owns NB consecutive blocks, window minima on 32-bit keys with an exact redo
// of the tiles that saw a tie. Covers the benchmark families with w = B + 1; sk4_supported() says which.
#ifndef SK4_THREADS
#define SK4_THREADS 128
#endif
#ifndef SK4_MIN_BLOCKS
#define SK4_MIN_BLOCKS 5
#endif
bool sk4_supported(int k, int m);
cudaError_t sk4_plan(SkArgs& args, int sm_count, SkLaunchPlan* plan);
// tickets one launch draws from the counter: one per tile plus one per warp (the draw that tells a warp to stop)
unsigned int sk4_ticket_draws(const SkArgs& args, const SkLaunchPlan& plan);
cudaError_t sk4_launch(const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream);

// sy4_kernel.cu: closed syncmers (blg_syncmers) for k - m = 12 or 10, k <= 32, in the same frame: sk4's 32-bit-key window minimum once per
// strand, keep bits from the packed argmin bytes. Uses sk4's scan scratch and ticket protocol (sk4_ticket_draws applies).
bool sy4_supported(int k, int m);
cudaError_t sy4_plan(SkArgs& args, int sm_count, SkLaunchPlan* plan);
cudaError_t sy4_launch(const SkArgs& args, const SkLaunchPlan& plan, cudaStream_t stream);

// kmer_view: every valid k-mer as (value, position), in position order.
struct KmerArgs {
    const unsigned char* bases;
    long long buf_lo, buf_hi;
    const unsigned long long* seq_offsets;
    long long n_seq;
    long long win_lo, win_hi;
    int k;
    int canonical;
    int vec_ok;
    int nwords;
    int steps;                     // kmer2: tile size chosen by kmer2_plan
    // closed-syncmer filter (sampler::syncmer_sampler over hash::minimizer_position_extractor): keep a k-mer iff the
    // offset of its minimum-hash m-mer is sync_s or sync_e; k <= 32 only; sync_on = 0: every valid k-mer
    int sync_on, sync_m, sync_s, sync_e;
    // sync_on = 2: sampler::hash_sampler instead (include/hash_sampler.hpp:138-141): keep iff hash64(value, seed) < samp_threshold
    unsigned long long samp_s8, samp_threshold;
    long long num_tiles;
    unsigned long long* value_lo;
    unsigned long long* value_hi;  // null for k <= 32
    unsigned long long* position;
    unsigned long long capacity;
    unsigned int* tile_counter;
    unsigned long long* tile_status;
    unsigned long long* total;
};
struct KmerLaunchPlan {
    int threads;
    size_t smem_bytes;
    int grid;
};
// kmer2_kernel.cu
cudaError_t kmer2_plan(KmerArgs& args, int sm_count, KmerLaunchPlan* plan);
cudaError_t kmer2_launch(const KmerArgs& args, const KmerLaunchPlan& plan, cudaStream_t stream);

// minimizer_context_t::id for records already on the device: rank directory over the valid-m-mer bit vector
size_t rank_scratch_bytes(long long n_bases);
cudaError_t record_ids_launch(const unsigned char* bases, long long n_bases, const unsigned long long* offs, long long n_seq,
                              int m, const unsigned long long* mm_start, unsigned long long n_records, void* scratch,
                              unsigned long long* id, cudaStream_t stream, int* launches);

// hash::minimizer_position_extractor for k-mer values held in uint64_t (k <= 32)
cudaError_t minpos_launch(const unsigned long long* values, size_t n, int k, int m, unsigned char* out, int sm_count, cudaStream_t stream);

// ingest_kernel.cu: 2-bit packed input -> ASCII; FASTA / FASTQ text -> bases + seq_offsets
cudaError_t unpack_range_launch(const void* codes, size_t word_lo, size_t word_hi, size_t n_bases, const unsigned long long* break_start,
                                const unsigned int* break_len, size_t n_breaks, unsigned char* out, cudaStream_t stream);
cudaError_t unpack_launch(const void* codes, size_t n_bases, const unsigned long long* break_start, const unsigned int* break_len,
                          size_t n_breaks, unsigned char* out, cudaStream_t stream);
size_t fastx_scratch_bytes(size_t n_text);
cudaError_t fastx_count_launch(const unsigned char* text, size_t n_text, int fastq, void* scratch, unsigned long long* totals, cudaStream_t stream,
                               int* launches);
cudaError_t fastx_emit_launch(const unsigned char* text, size_t n_text, int fastq, void* scratch, unsigned char* bases, unsigned long long* offsets,
                              unsigned long long* first_off, cudaStream_t stream, int* launches);

// payload_kernel.cu: records -> bucket-ordered records + their bases as one 2-bit packed stream
size_t payload_scratch_bytes(unsigned long long n_records, unsigned int n_buckets);
cudaError_t payload_launch(const unsigned char* bases, unsigned long long n_bases, const unsigned long long* value, const unsigned long long* mm_start, const unsigned char* size,
                           const unsigned char* mm_pos, unsigned long long n, int k, unsigned long long s8, unsigned int n_buckets, void* scratch,
                           unsigned long long* o_value, unsigned long long* o_mm, unsigned char* o_size, unsigned char* o_pos,
                           unsigned long long* o_pstart, unsigned long long* bucket_offsets, unsigned long long* words,
                           unsigned long long* totals, cudaStream_t stream, int* launches);

// sort_kernel.cu: radix sort / adjacent unique of 64-bit k-mer values (the reference's sorted external-memory vector +
// ordered_unique_sampler), the K1 drop-last filter of kmer_view
size_t sort_scratch_bytes(unsigned long long n);
cudaError_t radix_sort_u64(unsigned long long* keys, unsigned long long* tmp, unsigned long long n, int bits, void* scratch, cudaStream_t stream,
                           int* launches);
cudaError_t unique_sorted_u64(const unsigned long long* keys, unsigned long long n, unsigned long long* flags, void* scratch, unsigned long long* out,
                              unsigned long long* d_total, cudaStream_t stream, int* launches);
cudaError_t drop_last_kmers(unsigned long long* values, const unsigned long long* pos, unsigned long long n, int k, const unsigned long long* offs,
                            unsigned long long n_seq, unsigned long long n_bases, unsigned long long* flags, void* scratch, unsigned long long* tmp,
                            unsigned long long* d_total, cudaStream_t stream, int* launches);

cudaError_t partition_by_digit_u64(const unsigned long long* keys, unsigned long long* out, unsigned long long n, int shift, void* scratch,
                                   unsigned long long* bin_start, cudaStream_t stream, int* launches);
cudaError_t intersect_count_u64(const unsigned long long* a, unsigned long long na, const unsigned long long* b, unsigned long long nb,
                                unsigned long long* d_count, cudaStream_t stream);
cudaError_t lower_bound_u64(const unsigned long long* keys, unsigned long long n, const unsigned long long* d_queries, unsigned long long nq,
                            unsigned long long* d_out, cudaStream_t stream);

// BLG_F_RUNS: the maximal runs of records in which every record starts where the previous one ended
cudaError_t tight_pack_launch(const unsigned long long* value, const unsigned char* size, const unsigned char* mm_pos, unsigned long long n,
                              unsigned int* lo32, unsigned char* hi8, unsigned char* size_pos, cudaStream_t stream);
size_t runs_scratch_bytes(unsigned long long n);
cudaError_t runs_launch(const unsigned long long* mm_start, const unsigned char* mm_pos, const unsigned char* size, unsigned long long n, void* scratch,
                        unsigned long long** first, unsigned long long** start, unsigned long long* d_total, cudaStream_t stream, int* launches);

cudaError_t hash64_launch(const unsigned long long* in, unsigned long long* out, size_t n, unsigned long long s8,
                          cudaStream_t stream);

}  // namespace blg
