// biolib_gpu.hpp -- C++17 host shim over the C ABI (biolib_gpu.h): the reference's view classes, batched on a B200.
//
// Same constructor surface, accessors and iterator value types as the reference (file:line in /root/reference):
//   biolib::gpu::kmer_view<KmerType, Iterator>(start, stop, k, canonical = false)            include/kmer_view.hpp:61
//       kmer_view_from_string / kmer_view_from_cstr                                          :86-94
//       value_type kmer_context_t<KmerType>{optional value; position; id}                    :18-23
//   biolib::gpu::minimizer_view<KmerType, MinimizerType, HashFunction, Iterator>(start, stop, k, m, seed, canonical = false)
//                                                                                            include/minimizer_view.hpp:88
//       minimizer_view_from_string / _from_cstr                                              :103-122
//       value_type minimizer_context_t{value; position; id}                                  :21-25
//   biolib::gpu::super_kmer_view<KmerType, MinimizerType, HashFunction>(contig, len, k, m, canonical = false [, seed = 0])
//                                                                                            include/super_kmer_view.hpp:48-49
//       value_type super_kmer_t{minimizer; mm_pos; size}                                     :20-24
//   biolib::gpu::minimizer_position_extractor(k, m)                                          include/kmer_view.hpp:250-283
// A caller switches by replacing `wrapper::` with `biolib::gpu::` and linking libbiolib_gpu.so.
//
// Differences that are deliberate:
//   * the view runs the whole sequence through ONE device call the first time cbegin() is taken and then iterates
//     host copies of the records; iterators borrow from the view (same lifetime rule as the reference);
//   * Iterator must walk a contiguous char range (gpu::char_iterator, std::string::const_iterator, char const*);
//   * kmer_view reproduces the reference's record stream exactly, quirks included (K1 dropped last k-mer, K3 one
//     valueless record per break, K4 ids), on inputs where the reference loops forever (K6) it stops cleanly;
//   * minimizer_view / super_kmer_view yield the records the reference's (non-functional) iterators intend:
//     decisions D1-D9; `position` keeps the reference's 1-based formula (D7);
//   * HashFunction must be hash::hash64 (the only hash of the reference's views; MurmurHash3_x64_128 low word);
//   * errors surface as std::invalid_argument / std::runtime_error; there is NO CPU fallback: without a B200 and
//     the shared library every view throws on cbegin().
#ifndef BIOLIB_GPU_HPP
#define BIOLIB_GPU_HPP

#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <iterator>
#include <limits>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#include "biolib_gpu.h"

namespace hash { class hash64; }  // tag only; the reference defines it in include/hash.hpp:45

namespace biolib {
namespace gpu {

// ------------------------------------------------------------------------------------------------ context
class context
{
    public:
        explicit context(int device = 0) : h(nullptr)
        {
            const int st = blg_create(device, &h);
            if (st != BLG_OK) throw std::runtime_error(std::string("[biolib::gpu] no usable B200 context: ") + blg_strerror(st));
        }
        ~context() { if (h) blg_destroy(h); }
        context(context const&) = delete;
        context& operator=(context const&) = delete;
        blg_ctx* get() const noexcept { return h; }
        void check(int st, char const* what) const
        {
            if (st == BLG_OK) return;
            std::string msg = std::string("[biolib::gpu] ") + what + ": " + blg_strerror(st) + " (" + blg_last_error(h) + ")";
            if (st == BLG_E_INVALID) throw std::invalid_argument(msg);
            throw std::runtime_error(msg);
        }
        // one context per host thread (a blg_ctx is not thread-safe); device from BIOLIB_GPU_DEVICE, default 0
        static context& thread_default()
        {
            thread_local std::unique_ptr<context> ctx;
            if (!ctx) {
                char const* e = std::getenv("BIOLIB_GPU_DEVICE");
                ctx.reset(new context(e ? std::atoi(e) : 0));
            }
            return *ctx;
        }
    private:
        blg_ctx* h;
};

// same interface as the reference's char_iterator (include/constants.hpp:85-109) plus base(): the raw pointer
class char_iterator
{
    public:
        using iterator_category = std::forward_iterator_tag;
        using difference_type   = std::ptrdiff_t;
        using value_type        = char;
        using pointer           = value_type const*;
        using reference         = value_type const&;
        char_iterator(char const* ptr) : internal(ptr) {}
        value_type operator*() const noexcept { return *internal; }
        char_iterator const& operator++() { ++internal; return *this; }
        char_iterator operator++(int) { auto res = *this; ++internal; return res; }
        char const* base() const noexcept { return internal; }
        friend bool operator==(char_iterator const& a, char_iterator const& b) { return a.internal == b.internal; }
        friend bool operator!=(char_iterator const& a, char_iterator const& b) { return !(a == b); }
    private:
        char const* internal;
};

namespace detail {

// Contiguous char ranges only: raw pointers, std::string / std::vector<char> iterators, and gpu::char_iterator below.
// (The reference's own char_iterator, include/constants.hpp:85-109, hides its pointer and returns chars by value, so
// it cannot hand a buffer to a device call; use the *_from_cstr / *_from_string factories, as the reference's tests do.)
template <class Iterator, class = void>
struct has_base : std::false_type {};
template <class Iterator>
struct has_base<Iterator, std::void_t<decltype(std::declval<Iterator const&>().base())>>
    : std::is_convertible<decltype(std::declval<Iterator const&>().base()), char const*> {};

template <class Iterator>
char const* raw_pointer(Iterator it)
{
    if constexpr (std::is_pointer<Iterator>::value) return it;
    else if constexpr (has_base<Iterator>::value) return it.base();
    else {
        static_assert(std::is_same<Iterator, std::string::const_iterator>::value || std::is_same<Iterator, std::string::iterator>::value ||
                      std::is_same<Iterator, std::vector<char>::const_iterator>::value || std::is_same<Iterator, std::vector<char>::iterator>::value,
                      "biolib::gpu views need a contiguous char range (char const*, std::string iterators, gpu::char_iterator)");
        return std::addressof(*it);
    }
}

template <class Iterator>
std::size_t range_length(Iterator a, Iterator b)
{
    return static_cast<std::size_t>(raw_pointer(b) - raw_pointer(a));
}

template <class Iterator>
char const* contiguous_begin(Iterator it, std::size_t len)
{
    static_assert(std::is_same<typename std::iterator_traits<Iterator>::value_type, char>::value, "Iterator::value_type must be char");
    return len ? raw_pointer(it) : nullptr;
}

template <class T, class Free>
struct out_guard {
    context& c; T& o; Free f;
    ~out_guard() { f(c.get(), &o); }
};

}  // namespace detail

// ------------------------------------------------------------------------------------------------ kmer_view
template <typename KmerType>
struct kmer_context_t {          // layout of wrapper::kmer_context_t (include/kmer_view.hpp:18-23)
    std::optional<KmerType> value;
    std::size_t position;
    std::size_t id;
};

template <typename KmerType, class Iterator>
class kmer_view
{
    static_assert(sizeof(KmerType) == 8 || sizeof(KmerType) == 16, "KmerType is uint64_t or unsigned __int128");
    public:
        using value_type = kmer_context_t<KmerType>;
        class const_iterator
        {
            public:
                using iterator_category = std::forward_iterator_tag;
                using difference_type = std::ptrdiff_t;
                using value_type = kmer_context_t<KmerType>;
                using pointer = value_type const*;
                using reference = value_type const&;
                const_iterator(kmer_view const* v, std::size_t i) : parent(v), idx(i) {}
                value_type operator*() const noexcept { return parent->records[idx]; }   // by value, as the reference (:40)
                const_iterator const& operator++() { ++idx; return *this; }
                const_iterator operator++(int) { auto r = *this; ++idx; return r; }
                friend bool operator==(const_iterator const& a, const_iterator const& b) { return a.parent == b.parent && a.idx == b.idx; }
                friend bool operator!=(const_iterator const& a, const_iterator const& b) { return !(a == b); }
            private:
                kmer_view const* parent;
                std::size_t idx;
        };

        kmer_view(Iterator start, Iterator stop, uint8_t k, bool canonical = false)
            : itr_start(start), itr_stop(stop), klen(k), canon(canonical), done(false) {}
        const_iterator cbegin() const { run(); return const_iterator(this, 0); }
        const_iterator cend() const { run(); return const_iterator(this, records.size()); }
        const_iterator begin() const { return cbegin(); }
        const_iterator end() const { return cend(); }
        uint8_t get_k() const noexcept { return klen; }

    private:
        Iterator itr_start, itr_stop;
        uint8_t klen;
        bool canon;
        mutable bool done;
        mutable std::vector<value_type> records;

        void run() const
        {
            if (done) return;
            const std::size_t n = detail::range_length(itr_start, itr_stop);
            char const* s = detail::contiguous_begin(itr_start, n);
            if (2u * klen > 8 * sizeof(KmerType)) throw std::invalid_argument("[biolib::gpu::kmer_view] k does not fit KmerType");
            context& c = context::thread_default();
            blg_params p{};
            p.k = klen; p.canonical = canon ? 1 : 0;
            blg_kmer_out o{};
            c.check(blg_kmers(c.get(), s, n, nullptr, 0, &p, &o), "blg_kmers");
            detail::out_guard<blg_kmer_out, void (*)(blg_ctx*, blg_kmer_out*)> g{c, o, blg_free_kmer_out};
            // Rebuild the reference's record stream from the positions of the valid k-mers:
            // consecutive positions form a run; each run but the last is followed by one valueless record at
            // (last k-mer of the run) + 1; the final k-mer of a last run that touches the end of input is dropped.
            records.clear();
            records.reserve(o.count + 16);
            for (uint64_t i = 0; i < o.count; ++i) {
                const bool run_ends = (i + 1 == o.count) || (o.position[i + 1] != o.position[i] + 1);
                const bool last_run = run_ends && (i + 1 == o.count);
                if (last_run && o.position[i] + klen == n) break;                                      // K1
                KmerType v = static_cast<KmerType>(o.value_lo[i]);
                if constexpr (sizeof(KmerType) == 16) { if (o.value_hi) v |= static_cast<KmerType>(o.value_hi[i]) << 64; }
                records.push_back(value_type{v, static_cast<std::size_t>(o.position[i]), records.size()});
                if (run_ends && !last_run)
                    records.push_back(value_type{std::nullopt, static_cast<std::size_t>(o.position[i] + 1), records.size()});  // K3
            }
            done = true;
        }
};

template <typename KmerType>
kmer_view<KmerType, std::string::const_iterator> kmer_view_from_string(std::string const& s, uint8_t k, bool canonical)
{
    return kmer_view<KmerType, std::string::const_iterator>(s.cbegin(), s.cend(), k, canonical);
}

template <typename KmerType>
kmer_view<KmerType, char_iterator> kmer_view_from_cstr(char const* s, std::size_t len, uint8_t k, bool canonical)
{
    return kmer_view<KmerType, char_iterator>(char_iterator(s), char_iterator(s + len), k, canonical);
}

// ------------------------------------------------------------------------------------------------ minimizer_position_extractor
// hash::minimizer_position_extractor (reference include/kmer_view.hpp:250-283, ctor src/kmer_view.cpp:7-15): same
// constructor and accessors; operator() keeps the reference's one-record form (one device round trip per call: use it
// for spot checks) and gains a batched form over the records of a gpu::kmer_view or a plain array of k-mer values.
class minimizer_position_extractor
{
    public:
        using value_type = uint64_t;
        minimizer_position_extractor(uint8_t k, uint8_t m) : klen(k), mlen(m)
        {
            if (k == 0 || k > 32 || m == 0 || m > k) throw std::invalid_argument("[biolib::gpu] minimizer_position_extractor: need 1 <= m <= k <= 32");
        }
        uint8_t get_k() const noexcept { return klen; }
        uint8_t get_m() const noexcept { return mlen; }
        template <typename KmerType>
        std::size_t operator()(kmer_context_t<KmerType> const& kmer) const
        {
            static_assert(sizeof(KmerType) == 8, "k-mer values held in uint64_t");
            if (!kmer.value) return static_cast<std::size_t>(klen) + 1;          // include/kmer_view.hpp:270
            uint64_t v = *kmer.value;
            uint8_t out = 0;
            context& ctx = context::thread_default();
            ctx.check(blg_minimizer_positions(ctx.get(), &v, 1, klen, mlen, 0, &out), "minimizer_position_extractor");
            return out;
        }
        std::vector<uint8_t> operator()(uint64_t const* values, std::size_t n) const
        {
            std::vector<uint8_t> out(n);
            context& ctx = context::thread_default();
            ctx.check(blg_minimizer_positions(ctx.get(), values, n, klen, mlen, 0, out.data()), "minimizer_position_extractor");
            return out;
        }
    private:
        uint8_t klen, mlen;
};

// ------------------------------------------------------------------------------------------------ syncmer_sampler
// sampler::syncmer_sampler<kmer_view::const_iterator, hash::minimizer_position_extractor> (reference
// include/syncmer_sampler.hpp:9-66, selection rule :130-137), with the filter fused into the k-mer kernel (blg_syncmers):
// iterates the k-mers of [start, stop) whose extractor offset is start_offset or end_offset. The reference's sampler
// template also works unchanged over gpu::kmer_view iterators with gpu::minimizer_position_extractor (one device round
// trip per k-mer); this class is the batched form. Records are kmer_context_t with the k-mer's value and position;
// `id` here is the ordinal among the SAMPLED k-mers (the reference's is the kmer_view ordinal, which needs the unsampled
// stream). The final k-mer of an input that ends in a valid base is dropped, as the reference's iteration does.
template <class Iterator>
class syncmer_sampler
{
    public:
        using value_type = kmer_context_t<uint64_t>;
        using const_iterator = typename std::vector<value_type>::const_iterator;
        syncmer_sampler(Iterator start, Iterator stop, minimizer_position_extractor const& extractor, uint16_t start_offset,
                        uint16_t end_offset, bool canonical = false)
            : soffset(start_offset), eoffset(end_offset)
        {
            const std::size_t n = detail::range_length(start, stop);
            char const* s = detail::contiguous_begin(start, n);
            context& c = context::thread_default();
            blg_params p{};
            p.k = extractor.get_k(); p.m = extractor.get_m(); p.canonical = canonical ? 1 : 0;
            blg_kmer_out o{};
            c.check(blg_syncmers(c.get(), s, n, nullptr, 0, &p, start_offset, end_offset, &o), "blg_syncmers");
            detail::out_guard<blg_kmer_out, void (*)(blg_ctx*, blg_kmer_out*)> g{c, o, blg_free_kmer_out};
            uint64_t cnt = o.count;
            if (cnt > 0 && o.position[cnt - 1] + p.k == n) --cnt;                                       // K1
            records.reserve(cnt);
            for (uint64_t i = 0; i < cnt; ++i) records.push_back(value_type{o.value_lo[i], static_cast<std::size_t>(o.position[i]), static_cast<std::size_t>(i)});
        }
        const_iterator cbegin() const { return records.cbegin(); }
        const_iterator cend() const { return records.cend(); }
        const_iterator begin() const { return cbegin(); }
        const_iterator end() const { return cend(); }
        std::pair<uint16_t, uint16_t> get_offsets() const { return std::make_pair(soffset, eoffset); }
    private:
        uint16_t soffset, eoffset;
        std::vector<value_type> records;
};

// ------------------------------------------------------------------------------------------------ hash_sampler
// sampler::hash_sampler<Iterator, HashFunctionFamily>(start, stop, hash, seed, sampling_rate) (reference include/hash_sampler.hpp:73-78,
// selection rule :138-141: keep iff hash(value, seed) < threshold, threshold = (hash_type)(rate * max)) with the filter fused into
// the k-mer kernel (blg_hash_sampled_kmers). The reference's sampler walks an iterator over k-mer VALUES; here [start, stop) is
// the character range the k-mers come from, so k (and canonical) follow the reference's five arguments. The reference's own
// template also works unchanged over gpu::kmer_view iterators with hash::hash64. Records are kmer_context_t (value, position,
// id = ordinal among the sampled k-mers); the last k-mer of an input that ends in a valid base is dropped.
template <class Iterator, typename HashFunctionFamily>
class hash_sampler
{
    static_assert(std::is_same<HashFunctionFamily, ::hash::hash64>::value, "the device path implements hash::hash64 only");
    public:
        using value_type = kmer_context_t<uint64_t>;
        using const_iterator = typename std::vector<value_type>::const_iterator;
        hash_sampler(Iterator const& start, Iterator const& stop, HashFunctionFamily const& /*hash*/, uint64_t seed, double sampling_rate,
                     uint8_t k, bool canonical = false)
            : srate(sampling_rate)
        {
            if (srate > 1 or srate < 0) throw std::invalid_argument("[hash_sampler] Invalid sampling rate");       // hash_sampler.hpp:76
            // hash_sampler.hpp:77 computes static_cast<uint64_t>(srate * max); for rate 1.0 that cast is out of range (undefined
            // behaviour in the reference): the largest threshold is used there
            const double prod = srate * static_cast<double>(std::numeric_limits<uint64_t>::max());
            threshold = prod >= 18446744073709551615.0 ? std::numeric_limits<uint64_t>::max() : static_cast<uint64_t>(prod);
            const std::size_t n = detail::range_length(start, stop);
            char const* s = detail::contiguous_begin(start, n);
            context& c = context::thread_default();
            blg_params p{};
            p.k = k; p.canonical = canonical ? 1 : 0; p.seed = seed;
            blg_kmer_out o{};
            c.check(blg_hash_sampled_kmers(c.get(), s, n, nullptr, 0, &p, threshold, &o), "blg_hash_sampled_kmers");
            detail::out_guard<blg_kmer_out, void (*)(blg_ctx*, blg_kmer_out*)> g{c, o, blg_free_kmer_out};
            uint64_t cnt = o.count;
            if (cnt > 0 && o.position[cnt - 1] + k == n) --cnt;                                                       // K1
            records.reserve(cnt);
            for (uint64_t i = 0; i < cnt; ++i) records.push_back(value_type{o.value_lo[i], static_cast<std::size_t>(o.position[i]), static_cast<std::size_t>(i)});
        }
        const_iterator cbegin() const { return records.cbegin(); }
        const_iterator cend() const { return records.cend(); }
        const_iterator begin() const { return cbegin(); }
        const_iterator end() const { return cend(); }
        double get_sampling_rate() const { return srate; }
        uint64_t get_threshold() const { return threshold; }
    private:
        double srate;
        uint64_t threshold;
        std::vector<value_type> records;
};

// ------------------------------------------------------------------------------------------------ minimizer_view
template <typename MinimizerType>
struct minimizer_context_t {     // layout of minimizer_view::const_iterator::minimizer_context_t (minimizer_view.hpp:21-25)
    MinimizerType value;         // 2-bit packed minimizer (canonical-or-forward m-mer)
    std::size_t position;        // the reference's formula: 1-based start of the m-mer (D7)
    std::size_t id;              // ordinal of the m-mer among the valid m-mers of the sequence (D9)
};

namespace detail {

// shared by minimizer_view and super_kmer_view: ONE blg_super_kmers / blg_minimizers call; the library's own (pinned) result
// arrays are kept as they are -- structure of arrays -- and an iterator builds its record when it is dereferenced, so
// nothing is copied record by record. Inputs of 8 Mi k-mers and more take the library's chunked path (upload, kernel and
// record download overlap), ids included.
template <typename MinimizerType>
struct sk_records {
    static_assert(std::is_same<MinimizerType, uint64_t>::value, "MinimizerType must be uint64_t (m <= 32)");
    blg_sk_out o{};
    context* ctx = nullptr;

    sk_records() = default;
    sk_records(sk_records const&) = delete;
    sk_records& operator=(sk_records const&) = delete;
    ~sk_records() { if (ctx) blg_free_sk_out(ctx->get(), &o); }

    std::size_t size() const noexcept { return static_cast<std::size_t>(o.count); }
    minimizer_context_t<MinimizerType> minimizer(std::size_t i) const noexcept
    {
        return minimizer_context_t<MinimizerType>{o.value[i], static_cast<std::size_t>(o.mm_start[i] + 1),     // D7: 1-based start
                                                  static_cast<std::size_t>(o.id ? o.id[i] : i)};
    }
    // streaming form (super_kmer_view with want_ids = false): the smallest records the library has -- BLG_F_RUNS, and BLG_F_TIGHT
    // where (k, m) allow it -- no positions and no ids come back; a forward iterator rebuilds the positions by adding the sizes
    // inside a run, exactly what the reference's sequential view keeps implicitly
    bool streaming = false;
    bool tight() const noexcept { return o.size_pos != nullptr; }
    MinimizerType value_at(std::size_t i) const noexcept
    {
        if (!tight()) return o.value[i];
        return static_cast<MinimizerType>(o.value_lo32[i]) | (o.value_hi8 ? static_cast<MinimizerType>(o.value_hi8[i]) << 32 : 0);
    }
    uint8_t size_at(std::size_t i) const noexcept { return tight() ? static_cast<uint8_t>((o.size_pos[i] & 15u) + 1u) : o.size[i]; }
    uint8_t mm_pos_at(std::size_t i) const noexcept { return tight() ? static_cast<uint8_t>(o.size_pos[i] >> 4) : o.mm_pos[i]; }
    void run(char const* s, std::size_t n, uint8_t k, uint8_t m, uint64_t seed, bool canon, bool with_size, bool want_ids = true)
    {
        context& c = context::thread_default();
        blg_params p{};
        p.k = k; p.m = m; p.canonical = canon ? 1 : 0; p.seed = seed; p.flags = BLG_F_EMIT_ID;
        streaming = with_size && !want_ids;
        if (streaming) p.flags = BLG_F_RUNS | ((m <= 20 && k >= m && k - m <= 15) ? BLG_F_TIGHT : 0u);
        const int st = with_size ? blg_super_kmers(c.get(), s, n, nullptr, 0, &p, &o) : blg_minimizers(c.get(), s, n, nullptr, 0, &p, &o);
        ctx = &c;
        c.check(st, with_size ? "blg_super_kmers" : "blg_minimizers");
    }
};

}  // namespace detail

template <typename KmerType, typename MinimizerType, typename HashFunction, class Iterator>
class minimizer_view
{
    static_assert(std::is_same<HashFunction, ::hash::hash64>::value, "the device path implements hash::hash64 only");
    public:
        using value_type = minimizer_context_t<MinimizerType>;
        class const_iterator
        {
            public:
                using iterator_category = std::forward_iterator_tag;
                using difference_type = std::ptrdiff_t;
                using value_type = minimizer_context_t<MinimizerType>;
                using pointer = value_type const*;
                using reference = value_type const&;
                const_iterator(minimizer_view const* v, std::size_t i) : parent(v), idx(i) { load(); }
                value_type const& operator*() const noexcept { return current; }               // by reference (:34)
                const_iterator const& operator++() { ++idx; load(); return *this; }
                const_iterator operator++(int) { auto r = *this; ++idx; load(); return r; }
                friend bool operator==(const_iterator const& a, const_iterator const& b) { return a.parent == b.parent && a.idx == b.idx; }
                friend bool operator!=(const_iterator const& a, const_iterator const& b) { return !(a == b); }
            private:
                minimizer_view const* parent;
                std::size_t idx;
                value_type current{};
                void load() { if (idx < parent->rec.size()) current = parent->rec.minimizer(idx); }
        };

        minimizer_view(Iterator start, Iterator stop, uint8_t k, uint8_t m, uint64_t seed, bool canonical = false)
            : itr_start(start), itr_stop(stop), klen(k), mlen(m), mseed(seed), canon(canonical), done(false)
        {
            if (2u * klen > 8 * sizeof(KmerType)) throw std::invalid_argument("[biolib::gpu::minimizer_view] k does not fit KmerType");
        }
        const_iterator cbegin() const { run(); return const_iterator(this, 0); }
        const_iterator cend() const { run(); return const_iterator(this, rec.size()); }
        const_iterator begin() const { return cbegin(); }
        const_iterator end() const { return cend(); }
        uint8_t get_k() const noexcept { return klen; }
        uint8_t get_m() const noexcept { return mlen; }

    private:
        Iterator itr_start, itr_stop;
        uint8_t klen, mlen;
        uint64_t mseed;
        bool canon;
        mutable bool done;
        mutable detail::sk_records<MinimizerType> rec;
        void run() const
        {
            if (done) return;
            const std::size_t n = detail::range_length(itr_start, itr_stop);
            rec.run(detail::contiguous_begin(itr_start, n), n, klen, mlen, mseed, canon, false);
            done = true;
        }
};

template <typename KmerType, typename MinimizerType, typename HashFunction>
minimizer_view<KmerType, MinimizerType, HashFunction, std::string::const_iterator>
minimizer_view_from_string(std::string const& s, uint8_t k, uint8_t m, uint64_t seed, bool canonical)
{
    return minimizer_view<KmerType, MinimizerType, HashFunction, std::string::const_iterator>(s.cbegin(), s.cend(), k, m, seed, canonical);
}

template <typename KmerType, typename MinimizerType, typename HashFunction>
minimizer_view<KmerType, MinimizerType, HashFunction, char_iterator>
minimizer_view_from_cstr(char const* s, std::size_t len, uint8_t k, uint8_t m, uint64_t seed, bool canonical)
{
    return minimizer_view<KmerType, MinimizerType, HashFunction, char_iterator>(char_iterator(s), char_iterator(s + len), k, m, seed, canonical);
}

// ------------------------------------------------------------------------------------------------ super_kmer_view
template <typename MinimizerType>
struct super_kmer_t {            // super_kmer_view::const_iterator::super_kmer_t (super_kmer_view.hpp:20-24)
    minimizer_context_t<MinimizerType> minimizer;
    uint8_t mm_pos;              // position of the minimizer in the first k-mer
    uint8_t size;                // number of k-mers
};

template <typename KmerType, typename MinimizerType, typename HashFunction>
class super_kmer_view
{
    static_assert(std::is_same<HashFunction, ::hash::hash64>::value, "the device path implements hash::hash64 only");
    public:
        using value_type = super_kmer_t<MinimizerType>;
        class const_iterator
        {
            public:
                using iterator_category = std::forward_iterator_tag;
                using difference_type = std::ptrdiff_t;
                using value_type = super_kmer_t<MinimizerType>;
                using pointer = value_type const*;
                using reference = value_type const&;
                const_iterator(super_kmer_view const* v, std::size_t i) : parent(v), idx(i) { load(); }
                value_type const& operator*() const { return current_sk; }
                const_iterator const& operator++() { ++idx; load(); return *this; }
                const_iterator operator++(int) { auto r = *this; ++idx; load(); return r; }
                friend bool operator==(const_iterator const& a, const_iterator const& b) { return a.parent == b.parent && a.idx == b.idx; }
                friend bool operator!=(const_iterator const& a, const_iterator const& b) { return !(a == b); }
            private:
                super_kmer_view const* parent;
                std::size_t idx;
                value_type current_sk;
                std::size_t run = 0;          // streaming form: the run the record belongs to, and where its first k-mer starts
                uint64_t sk_start = 0;
                void load()
                {
                    auto const& r = parent->rec;
                    if (idx >= r.size()) return;
                    if (!r.streaming) { current_sk = value_type{r.minimizer(idx), r.o.mm_pos[idx], r.o.size[idx]}; return; }
                    // records are visited in order (forward iterator): a record starts where its predecessor's k-mers end, unless it
                    // opens the next run (a break or the start of the input lies between the two)
                    if (run < r.o.n_runs && r.o.run_first_record[run] == idx) sk_start = r.o.run_start[run++];
                    else sk_start += r.size_at(idx - 1);
                    const uint8_t mp = r.mm_pos_at(idx);
                    current_sk = value_type{minimizer_context_t<MinimizerType>{r.value_at(idx), static_cast<std::size_t>(sk_start + mp + 1), idx},
                                            mp, r.size_at(idx)};
                }
        };

        // the reference's constructors have no seed (super_kmer_view.hpp:48-49); it is added LAST with the default 0.
        // want_ids = false selects the streaming form: minimizer.id then holds the record's ordinal instead of SURVEY D9's m-mer
        // count, iterators must be advanced from cbegin() (they are forward iterators anyway), and the call moves 6 or 10 bytes per
        // record over PCIe instead of 26 (the id pass over the bases is skipped as well)
        super_kmer_view(char const* contig, std::size_t contig_len, uint8_t k, uint8_t m, bool canonical = false, uint64_t seed = 0,
                        bool want_ids = true)
            : seq(contig), slen(contig_len), klen(k), mlen(m), mseed(seed), canon(canonical), ids(want_ids), done(false)
        {
            if (!contig && contig_len) throw std::runtime_error("[super k-mer view] null contig");
            if (2u * klen > 8 * sizeof(KmerType)) throw std::invalid_argument("[biolib::gpu::super_kmer_view] k does not fit KmerType");
        }
        super_kmer_view(std::string const& contig, uint8_t k, uint8_t m, bool canonical = false, uint64_t seed = 0, bool want_ids = true)
            : super_kmer_view(contig.data(), contig.size(), k, m, canonical, seed, want_ids) {}
        const_iterator cbegin() const { run(); return const_iterator(this, 0); }
        const_iterator cend() const { run(); return const_iterator(this, rec.size()); }
        const_iterator begin() const { return cbegin(); }
        const_iterator end() const { return cend(); }
        uint8_t get_k() const noexcept { return klen; }
        uint8_t get_m() const noexcept { return mlen; }

    private:
        char const* seq;
        std::size_t slen;
        uint8_t klen, mlen;
        uint64_t mseed;
        bool canon, ids;
        mutable bool done;
        mutable detail::sk_records<MinimizerType> rec;
        void run() const
        {
            if (done) return;
            rec.run(seq, slen, klen, mlen, mseed, canon, true, ids);
            done = true;
        }
};

}  // namespace gpu
}  // namespace biolib

#endif  // BIOLIB_GPU_HPP
