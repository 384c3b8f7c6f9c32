// TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path.
//
// Tier A oracle: the reference's own code, executed verbatim.
//
// This translation unit contains NO algorithm of its own. It #includes the reference's
// headers where they lie (-I/root/reference/include, -I/root/reference/bundled; the
// recipe is oracle/Makefile, outputs go to oracle/_ref/ only) and exposes a C ABI over
//   * wrapper::kmer_view iteration                      (include/kmer_view.hpp:155-234)
//   * wrapper::minimizer_view::const_iterator::read_base (include/minimizer_view.hpp:229-245)
//   * sampler::syncmer_sampler over hash::minimizer_position_extractor (include/syncmer_sampler.hpp:130-137)
//   * hash::hash64 / hash::double_hash64                (include/hash.hpp:11-71)
//   * sequence::generator, wrapper::sequence_mutator    (include/sequence_generator.hpp:14-43,
//                                                        include/sequence_mutator.hpp:109-169)
// so the C restatement in oracle/oracle.c (and, through it, the CUDA path) can be pinned
// against reference-executed results. Guards applied before calling into the reference
// (SURVEY.md section 8(a), K5/K6/K8):
//   K5  iterators are placement-constructed in zero-filled storage (kmer_buffer is never
//       initialised by the reference's constructor); this TU must be compiled with
//       -fno-lifetime-dse (oracle/Makefile), otherwise GCC deletes the zero-fill as a dead
//       store before the constructor and canonical k-mers come out of stack garbage;
//   K6  the input handed to kmer_view is truncated right after the first invalid byte
//       that follows the last run of >= k valid bases (the reference over-reads and never
//       terminates otherwise); no such run => the reference is not called at all;
//   K8  bytes >= 0x80 are replaced by 'N' (negative char index is UB / throws).
//
// minimizer_view.hpp is not self-contained (needs <vector>, <cassert>) and read_base()
// is private: all standard headers are included first, then `private` is opened for the
// reference headers only.

#include <array>
#include <cassert>
#include <cstdint>
#include <cstring>
#include <limits>
#include <new>
#include <optional>
#include <random>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#define private public
#include "constants.hpp"
#include "hash.hpp"
#include "kmer_view.hpp"
#include "minimizer_view.hpp"
#include "syncmer_sampler.hpp"
#include "hash_sampler.hpp"
#undef private
#include "sequence_generator.hpp"
#include "sequence_mutator.hpp"
#include "packed_vector.hpp"
#include <mutex>
#include <tuple>
#include "ordered_unique_sampler.hpp"
#include "jaccard.hpp"

// kseq.h (reference tests/kseq.h, the parser tests/test_kmer_view.cpp:30-40 reads its input with) over an in-memory reader
struct ref_membuf { const char* p; size_t n, pos; };
static int ref_memread(ref_membuf* b, void* dst, int want)
{
    const size_t left = b->n - b->pos, take = left < (size_t)want ? left : (size_t)want;
    memcpy(dst, b->p + b->pos, take);
    b->pos += take;
    return (int)take;
}
#include "kseq.h"
KSEQ_INIT(ref_membuf*, ref_memread)

namespace {

inline bool is_valid_base(unsigned char c) { return constants::seq_nt4_table[c] < 4; }

// K6 + K8: sanitised, truncated copy that is inside kmer_view's defined domain.
// Returns false when the reference must not be called (no run of >= k valid bases).
bool kmer_view_safe_input(const char* s, size_t n, unsigned k, std::string& out)
{
    out.assign(s, n);
    for (auto& ch : out)
        if (static_cast<unsigned char>(ch) >= 0x80) ch = 'N';
    size_t run = 0, last_good_end = 0;  // end (exclusive) of the last run of >= k valid bases
    bool found = false;
    for (size_t i = 0; i <= n; ++i) {
        bool v = i < n && is_valid_base(static_cast<unsigned char>(out[i]));
        if (v) { ++run; continue; }
        if (run >= k) { found = true; last_good_end = i; }
        run = 0;
    }
    if (!found) return false;
    // keep [0, last_good_end) plus at most one following (invalid) byte
    if (last_good_end < n) out.resize(last_good_end + 1);
    else out.resize(last_good_end);
    return true;
}

template <typename KmerT>
size_t run_kmer_view(const char* s, size_t n, unsigned k, int canonical,
                     KmerT* value, uint8_t* has_value, uint64_t* position, uint64_t* id, size_t cap)
{
    std::string safe;
    if (k == 0 || !kmer_view_safe_input(s, n, k, safe)) return 0;
    auto view = wrapper::kmer_view_from_cstr<KmerT>(safe.data(), safe.size(), static_cast<uint8_t>(k), canonical != 0);
    using view_t = decltype(view);
    using it_t = typename view_t::const_iterator;
    alignas(it_t) unsigned char storage[sizeof(it_t)];
    std::memset(storage, 0, sizeof(storage));               // K5
    it_t* it = new (storage) it_t(&view);
    auto end = view.cend();
    size_t cnt = 0;
    for (; *it != end; ++(*it)) {
        auto rec = **it;
        if (cnt < cap) {
            if (value) value[cnt] = rec.value ? *rec.value : KmerT(0);
            if (has_value) has_value[cnt] = rec.value ? 1 : 0;
            if (position) position[cnt] = rec.position;
            if (id) id[cnt] = rec.id;
        }
        ++cnt;
    }
    it->~it_t();
    return cnt;
}

} // namespace

// Per-base stream of minimizer_view::const_iterator::read_base() + HashFunction::hash,
// driven exactly as init_window()/operator++ drive it (minimizer_view.hpp:194-203,
// 252-261) but without the (non-functional, F2) window. For every input byte i:
//   idx[i]   = value returned by read_base()  (= 1-based m-mer start, D7)
//   bslb[i]  = bases_since_last_break after the call
//   valid[i] = bslb[i] >= m ; value[i] = mm_forward_reverse[strand] ; hash[i] = hash64(value, seed)
// KmerType is __uint128_t when k > 32 (as tests/test_super_kmer_view.cpp:17 does), else uint64_t;
// it only feeds the mask computation (minimizer_view.hpp:172).
template <typename KmerT>
static void mmer_stream_impl(const char* s, size_t n, unsigned k, unsigned m, uint64_t seed, int canonical,
                             uint64_t* value, uint64_t* idx, uint64_t* hashv, uint64_t* bslb, uint8_t* valid)
{
    using view_t = wrapper::minimizer_view<KmerT, uint64_t, hash::hash64, char_iterator>;
    view_t view(char_iterator(s), char_iterator(s + n), static_cast<uint8_t>(k), static_cast<uint8_t>(m), seed, canonical != 0);
    typename view_t::const_iterator it(&view, 0);          // end-constructor: no init_window()
    it.itr = char_iterator(s);
    it.shift = 2 * (m - 1);                                // as minimizer_view.hpp:171-173
    it.mask = (KmerT(1) << (2 * m)) - 1;
    it.mm_forward_reverse = {uint64_t(0), uint64_t(0)};
    for (size_t i = 0; i < n; ++i) {
        auto r = it.read_base();
        bool ok = it.bases_since_last_break >= m;
        uint64_t v = it.mm_forward_reverse[it.strand];
        if (idx) idx[i] = r;
        if (bslb) bslb[i] = it.bases_since_last_break;
        if (valid) valid[i] = ok;
        if (value) value[i] = ok ? v : 0;
        if (hashv) hashv[i] = ok ? hash::hash64::hash(v, view.mseed) : 0;
    }
}

extern "C" {

// nt4 table as the reference holds it (include/constants.hpp:12-21)
void ref_nt4_table(uint8_t out[256]) { for (int i = 0; i < 256; ++i) out[i] = constants::seq_nt4_table[i]; }

uint64_t ref_hash64_u64(uint64_t v, uint64_t seed) { return hash::hash64::hash(v, seed); }
void ref_double_hash64_u64(uint64_t v, uint64_t seed, uint64_t out[2]) { auto h = hash::double_hash64::hash(v, seed); out[0] = h[0]; out[1] = h[1]; }
void ref_double_hash64_u128(uint64_t lo, uint64_t hi, uint64_t seed, uint64_t out[2])
{
    __uint128_t v = (static_cast<__uint128_t>(hi) << 64) | lo;
    auto h = hash::double_hash64::hash(v, seed);
    out[0] = h[0]; out[1] = h[1];
}
uint64_t ref_remix(uint64_t z) { return hash::remix(z); }

// hash::minimizer_position_extractor::operator() verbatim (include/kmer_view.hpp:266-283, ctor src/kmer_view.cpp:7-15),
// for k-mers held in uint64_t (k <= 32). has_value == 0 exercises the nullopt branch (returns k + 1).
void ref_minimizer_positions(const uint64_t* values, const uint8_t* has_value, size_t n, unsigned k, unsigned m, uint64_t* out)
{
    hash::minimizer_position_extractor ex(static_cast<uint8_t>(k), static_cast<uint8_t>(m));
    for (size_t i = 0; i < n; ++i) {
        wrapper::kmer_context_t<uint64_t> ctx;
        if (!has_value || has_value[i]) ctx.value = values[i]; else ctx.value = std::nullopt;
        ctx.position = i; ctx.id = i;
        out[i] = ex(ctx);
    }
}

// Full kmer_view stream, including nullopt records and ids (K1-K4).
size_t ref_kmer_view_u64(const char* s, size_t n, unsigned k, int canonical,
                         uint64_t* value, uint8_t* has_value, uint64_t* position, uint64_t* id, size_t cap)
{
    return run_kmer_view<uint64_t>(s, n, k, canonical, value, has_value, position, id, cap);
}

// sampler::syncmer_sampler<kmer_view<uint64_t>::const_iterator, hash::minimizer_position_extractor> driven over the
// reference's kmer_view (same guards as run_kmer_view). operator* of the sampler does not instantiate for this pair
// (kmer_context_t does not convert to value_type), so the current record is read through the (opened) itr_start member.
size_t ref_syncmers_u64(const char* s, size_t n, unsigned k, unsigned m, unsigned soff, unsigned eoff, int canonical,
                        uint64_t* value, uint64_t* position, uint64_t* id, size_t cap)
{
    std::string safe;
    if (k == 0 || !kmer_view_safe_input(s, n, k, safe)) return 0;
    auto view = wrapper::kmer_view_from_cstr<uint64_t>(safe.data(), safe.size(), static_cast<uint8_t>(k), canonical != 0);
    using it_t = decltype(view)::const_iterator;
    alignas(it_t) unsigned char storage[sizeof(it_t)];
    std::memset(storage, 0, sizeof(storage));               // K5
    it_t* first = new (storage) it_t(&view);
    hash::minimizer_position_extractor ex(static_cast<uint8_t>(k), static_cast<uint8_t>(m));
    sampler::syncmer_sampler<it_t, hash::minimizer_position_extractor> smp(*first, view.cend(), ex, static_cast<uint16_t>(soff), static_cast<uint16_t>(eoff));
    size_t cnt = 0;
    for (auto it = smp.cbegin(); it != smp.cend(); ++it) {
        auto rec = *it.itr_start;
        if (cnt < cap) {
            value[cnt] = rec.value ? *rec.value : 0;
            position[cnt] = rec.position;
            id[cnt] = rec.id;
        }
        ++cnt;
    }
    first->~it_t();
    return cnt;
}

// sampler::hash_sampler<std::vector<uint64_t>::const_iterator, hash::hash64> (include/hash_sampler.hpp:73-78,138-141)
// over an array of k-mer values: the kept values, and the threshold the constructor derived from the rate.
size_t ref_hash_sampler_u64(const uint64_t* values, size_t n, uint64_t seed, double rate, uint64_t* kept, uint64_t* threshold)
{
    std::vector<uint64_t> v(values, values + n);
    using it_t = std::vector<uint64_t>::const_iterator;
    sampler::hash_sampler<it_t, hash::hash64> smp(v.cbegin(), v.cend(), hash::hash64(), seed, rate);
    if (threshold) *threshold = smp.threshold;
    size_t cnt = 0;
    for (auto it = smp.cbegin(); it != smp.cend(); ++it) kept[cnt++] = *it;
    return cnt;
}

// 128-bit variant: values returned as (lo, hi) pairs. Canonical output for k > 32 is the
// reference's (wrong, K7) result and is only used to document the divergence.
size_t ref_kmer_view_u128(const char* s, size_t n, unsigned k, int canonical,
                          uint64_t* value_lo_hi, uint8_t* has_value, uint64_t* position, uint64_t* id, size_t cap)
{
    std::vector<__uint128_t> tmp(cap);
    size_t cnt = run_kmer_view<__uint128_t>(s, n, k, canonical, tmp.data(), has_value, position, id, cap);
    for (size_t i = 0; i < cnt && i < cap; ++i) {
        value_lo_hi[2 * i] = static_cast<uint64_t>(tmp[i]);
        value_lo_hi[2 * i + 1] = static_cast<uint64_t>(tmp[i] >> 64);
    }
    return cnt;
}

void ref_mmer_stream(const char* s, size_t n, unsigned k, unsigned m, uint64_t seed, int canonical,
                     uint64_t* value, uint64_t* idx, uint64_t* hashv, uint64_t* bslb, uint8_t* valid)
{
    if (k > 32) mmer_stream_impl<__uint128_t>(s, n, k, m, seed, canonical, value, idx, hashv, bslb, valid);
    else mmer_stream_impl<uint64_t>(s, n, k, m, seed, canonical, value, idx, hashv, bslb, valid);
}

// The literal minimizer_view iteration (F2: yields nothing). Returns the number of records.
size_t ref_minimizer_view_literal_count(const char* s, size_t n, unsigned k, unsigned m, uint64_t seed, int canonical)
{
    auto view = wrapper::minimizer_view_from_cstr<uint64_t, uint64_t, hash::hash64>(s, n, static_cast<uint8_t>(k), static_cast<uint8_t>(m), seed, canonical != 0);
    size_t cnt = 0;
    for (auto it = view.cbegin(); it != view.cend(); ++it) ++cnt;
    return cnt;
}

// sequence::generator::nucleic<packed::uniform>::get_sequence
void ref_generate(uint64_t seed, size_t len, char* out)
{
    sequence::generator::packed::uniform gen(seed);
    sequence::generator::nucleic<sequence::generator::packed::uniform> nuc(gen);
    std::string s = nuc.get_sequence(len);
    std::memcpy(out, s.data(), len);
}

// wrapper::sequence_mutator driven as tests/test_random_sequence_generation.cpp:44-48 does
// (nullopt dropped). out must hold worst-case growth; returns the mutated length.
// counts = {substitutions, insertions, deletions}.
size_t ref_mutate(const char* s, size_t n, double mutp, double indelf, double extp, uint64_t seed,
                  char* out, size_t cap, uint64_t counts[3], int64_t* total_indel_len)
{
    std::string src(s, n);
    wrapper::sequence_mutator<std::string::const_iterator> mut(src.cbegin(), src.cend(), mutp, indelf, extp);
    size_t o = 0;
    auto it = mut.cbegin(seed);
    for (; it != mut.cend(); ++it) {
        auto v = *it;
        if (v) { if (o < cap) out[o] = *v; ++o; }
    }
    auto rep = it.get_report();
    if (counts) { counts[0] = rep.substitution_count; counts[1] = rep.insertion_count; counts[2] = rep.deletion_count; }
    if (total_indel_len) *total_indel_len = rep.total_indel_len;
    return o;
}

// sampler::ordered_unique_sampler over two SORTED value lists + algorithm::jaccard, verbatim (include/ordered_unique_sampler.hpp:116-130,
// include/jaccard.hpp:8-38; tests/test_jaccard.cpp:97-130 is the same loop): out = {intersection, union, size1, size2}
void ref_jaccard_sorted(const uint64_t* a, size_t na, const uint64_t* b, size_t nb, uint64_t out[4])
{
    const std::vector<uint64_t> va(a, a + na), vb(b, b + nb);      // the sampler wants class-type iterators (Iterator::value_type)
    sampler::ordered_unique_sampler ua(va.cbegin(), va.cend());
    sampler::ordered_unique_sampler ub(vb.cbegin(), vb.cend());
    auto [inter, uni, s1, s2] = algorithm::jaccard(ua.cbegin(), ua.cend(), ub.cbegin(), ub.cend());
    out[0] = inter; out[1] = uni; out[2] = s1; out[3] = s2;
}

// the distinct values of a sorted list as ordered_unique_sampler yields them
size_t ref_ordered_unique(const uint64_t* a, size_t na, uint64_t* out)
{
    const std::vector<uint64_t> va(a, a + na);
    sampler::ordered_unique_sampler ua(va.cbegin(), va.cend());
    size_t n = 0;
    for (auto it = ua.cbegin(); it != ua.cend(); ++it) out[n++] = *it;
    return n;
}

// bit::packed::vector<uint64_t>(bitwidth): push_back every value, return the underlying words (include/packed_vector.hpp:134-160)
size_t ref_packed_vector_words(const uint8_t* values, size_t n, unsigned bitwidth, uint64_t* out_words, size_t out_cap)
{
    bit::packed::vector<uint64_t> pv(bitwidth);
    for (size_t i = 0; i < n; ++i) pv.push_back(values[i]);
    const size_t nw = pv.underlying_size();
    for (size_t i = 0; i < nw && i < out_cap; ++i) out_words[i] = pv.data()[i];
    return nw;
}

// kseq_read() until the end of the text: the sequences concatenated + their offsets (n_seq + 1 entries). Returns n_seq.
size_t ref_kseq_parse(const char* text, size_t n, char* bases_out, uint64_t* offs_out)
{
    ref_membuf mb{text, n, 0};
    kseq_t* seq = kseq_init(&mb);
    size_t n_seq = 0;
    uint64_t total = 0;
    offs_out[0] = 0;
    while (kseq_read(seq) >= 0) {
        memcpy(bases_out + total, seq->seq.s, seq->seq.l);
        total += seq->seq.l;
        offs_out[++n_seq] = total;
    }
    kseq_destroy(seq);
    return n_seq;
}

} // extern "C"

// CPU arm of the benchmark: the reference's own per-base arithmetic executed verbatim
// (read_base(): minimizer_view.hpp:229-245; hash::hash64::hash: hash.hpp:55-59 -> bundled Murmur)
// feeding the restated window-minimum / cut rules D4-D8 (monotone deque, as oracle.c mode 1) -- the
// reference's own window never fills (F2), so this is the closest runnable "reference CPU path".
// Single sequence. Returns the record count; sums[0..3] = wrapping sums of value, mm_start, size, hash.
template <typename KmerT>
static size_t super_kmers_sum_impl(const char* s, size_t n, unsigned k, unsigned m, uint64_t seed, int canonical, uint64_t sums[4])
{
    using view_t = wrapper::minimizer_view<KmerT, uint64_t, hash::hash64, char_iterator>;
    view_t view(char_iterator(s), char_iterator(s + n), static_cast<uint8_t>(k), static_cast<uint8_t>(m), seed, canonical != 0);
    typename view_t::const_iterator it(&view, 0);
    it.itr = char_iterator(s);
    it.shift = 2 * (m - 1);
    it.mask = (KmerT(1) << (2 * m)) - 1;
    it.mm_forward_reverse = {uint64_t(0), uint64_t(0)};
    const unsigned w = k - m + 1;
    uint64_t hring[128], vring[128];   // 128: with w = 64 m-mer cur_arg + 64 arrives before cur_arg's record closes
    size_t dq[128];
    unsigned head = 0, len = 0;
    size_t cnt = 0, cur_arg = 0, cur_n = 0;
    bool open = false;
    uint64_t sv = 0, sp = 0, ss = 0, sh = 0;
    auto emit = [&]() { sv += vring[cur_arg & 127]; sp += cur_arg; ss += cur_n; sh += hring[cur_arg & 127]; ++cnt; open = false; };
    for (size_t i = 0; i < n; ++i) {
        it.read_base();
        const size_t b = it.bases_since_last_break;
        if (b == 0) { if (open) emit(); len = 0; continue; }
        if (b < m) continue;
        const size_t j = i + 1 - m;
        const uint64_t v = it.mm_forward_reverse[it.strand];
        const uint64_t h = hash::hash64::hash(v, view.mseed);
        hring[j & 127] = h; vring[j & 127] = v;
        while (len && hring[dq[(head + len - 1) & 127] & 127] > h) --len;
        dq[(head + len) & 127] = j; ++len;
        if (b < k) continue;
        const size_t st = i + 1 - k;
        while (dq[head & 127] < st) { head = (head + 1) & 127; --len; }
        const size_t arg = dq[head & 127];
        if (open && arg == cur_arg) ++cur_n;
        else { if (open) emit(); open = true; cur_arg = arg; cur_n = 1; }
    }
    if (open) emit();
    (void)w;
    if (sums) { sums[0] = sv; sums[1] = sp; sums[2] = ss; sums[3] = sh; }
    return cnt;
}

extern "C" size_t ref_super_kmers_sums(const char* s, size_t n, unsigned k, unsigned m, uint64_t seed, int canonical, uint64_t sums[4])
{
    if (m == 0 || m > 32 || k < m || k > 64) return 0;
    if (k > 32) return super_kmers_sum_impl<__uint128_t>(s, n, k, m, seed, canonical, sums);
    return super_kmers_sum_impl<uint64_t>(s, n, k, m, seed, canonical, sums);
}
