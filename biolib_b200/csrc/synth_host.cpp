// synth_host.cpp -- host-side helpers of the C ABI (include/biolib_gpu.h): the workload generators blg_synth_generate /
// blg_synth_mutate and the 2-bit packer blg_pack_bases (the wire format of blg_super_kmers_packed). They restate, for batched buffers, the two generators the reference's benchmarks draw their inputs
// from:
//   sequence::generator::nucleic<packed::uniform>::get_sequence   reference include/sequence_generator.hpp:14-23,28-43
//   wrapper::sequence_mutator<Iterator>::const_iterator           reference include/sequence_mutator.hpp:109-169
// Both are defined by libstdc++'s std::mt19937 and distributions, which are used here directly (same library, same
// draws). They run on the host like the reference's own; nothing on the hot path calls them.
#include <cstddef>
#include <cstdint>
#include <random>

#include "../../include/biolib_gpu.h"

namespace {

const char kBases[4] = {'A', 'C', 'G', 'T'};     // constants::bases, reference include/constants.hpp:23

inline int code_of(char ch)                       // constants::seq_nt4_table restricted to its valid entries
{
    switch (ch) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': case 'U': case 'u': return 3;
    default: return 4;
    }
}

}  // namespace

extern "C" {

// ASCII -> 16 bases per 32-bit word (base q at bits [2q, 2q+2)) + the intervals of bytes that are not AaCcGgTtUu
// (constants::seq_nt4_table, reference include/constants.hpp:12-21). A code under a break is 0.
int blg_pack_bases(const char* bases, size_t n, uint32_t* codes, uint64_t* break_start, uint32_t* break_len, size_t break_capacity,
                   size_t* n_breaks)
{
    if ((n && (!bases || !codes)) || !n_breaks) return BLG_E_INVALID;
    size_t nb = 0;
    bool in_break = false;
    uint64_t run_start = 0;
    auto close_run = [&](uint64_t end) {
        // intervals longer than 32 bits of length are split
        uint64_t s0 = run_start;
        while (s0 < end) {
            const uint64_t len = end - s0 > 0xFFFFFFFFull ? 0xFFFFFFFFull : end - s0;
            if (nb < break_capacity) { break_start[nb] = s0; break_len[nb] = (uint32_t)len; }
            ++nb;
            s0 += len;
        }
    };
    const size_t words = (n + 15) / 16;
    for (size_t w = 0; w < words; ++w) {
        uint32_t word = 0;
        const size_t lo = w * 16, hi = lo + 16 < n ? lo + 16 : n;
        for (size_t i = lo; i < hi; ++i) {
            const int c = code_of(bases[i]);
            if (c > 3) { if (!in_break) { in_break = true; run_start = i; } }
            else {
                if (in_break) { in_break = false; close_run(i); }
                word |= (uint32_t)c << (2 * (i - lo));
            }
        }
        codes[w] = word;
    }
    if (in_break) close_run(n);
    *n_breaks = nb;
    return nb > break_capacity ? BLG_E_CAPACITY : BLG_OK;
}

int blg_synth_generate(uint64_t seed, size_t length, char* out)
{
    if (length && !out) return BLG_E_INVALID;
    std::mt19937 engine(static_cast<std::mt19937::result_type>(seed));   // the seed narrows to 32 bits, as in the reference
    std::uniform_int_distribution<uint8_t> dist(0, 3);
    for (size_t i = 0; i < length; ++i) out[i] = kBases[dist(engine)];
    return BLG_OK;
}

// The mutator's output stream with the absent values dropped (the way tests/test_random_sequence_generation.cpp:44-48
// consumes it). Behaviours kept because they are what the reference yields: the last source base is never emitted (the
// end test fires right after it is read); a deletion of length L consumes ONE source base; an insertion repeats the
// base L extra times; mutations never produce N. A source byte outside AaCcGgTtUu is BLG_E_INVALID (the reference
// indexes bases.at(4) there).
int blg_synth_mutate(const char* src, size_t n, double mutp, double indelf, double extp, uint64_t seed, char* out,
                     size_t capacity, size_t* out_len, uint64_t counts[3], int64_t* total_indel_len)
{
    if ((n && !src) || !out_len) return BLG_E_INVALID;
    std::mt19937 engine(static_cast<std::mt19937::result_type>(seed));
    std::uniform_real_distribution<double> probability(0.0, 1.0);
    size_t produced = 0, next = 0;
    uint64_t subs = 0, ins = 0, dels = 0;
    int64_t indel_total = 0;
    // one step of the iterator's operator++: sets (present, value) and consumes a source base unless an indel is running
    bool present = false;
    char value = 0;
    size_t pending = 0;       // further steps that repeat the current output (indel_len of the reference)
    auto step = [&]() -> int {
        if (pending) { --pending; return BLG_OK; }
        const char base = src[next++];
        const int c = code_of(base);
        if (c > 3) return BLG_E_INVALID;
        present = true;
        value = base;
        if (probability(engine) < mutp) {
            if (probability(engine) >= indelf) {
                const int shifted = (c + static_cast<uint8_t>(probability(engine) * 3 + 1)) & 3;
                ++subs;
                value = kBases[shifted];
            } else {
                size_t len = 1;
                while (probability(engine) < extp) ++len;
                pending = len;
                if (probability(engine) < 0.5) { ++dels; indel_total -= (int64_t)len; present = false; }
                else { ++ins; indel_total += (int64_t)len; value = kBases[c]; }
            }
        }
        return BLG_OK;
    };
    if (n) {
        int rc = step();                            // the begin iterator's constructor advances once
        if (rc != BLG_OK) return rc;
        while (next != n) {                         // iterator != cend(): compares the source position only
            if (present) { if (produced < capacity) out[produced] = value; ++produced; }
            if ((rc = step()) != BLG_OK) return rc;
        }
    }
    *out_len = produced;
    if (counts) { counts[0] = subs; counts[1] = ins; counts[2] = dels; }
    if (total_indel_len) *total_indel_len = indel_total;
    return produced > capacity ? BLG_E_CAPACITY : BLG_OK;
}

}  // extern "C"
