/*
 * TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and there only as the checker / the CPU arm.
 *
 * Tier B oracle: a plain-C CPU restatement of the reference's streaming k-mer path
 * (yhhshb/biolib). Citations are file:line relative to /root/reference.
 *
 * Parity status: PINNED for nt4, hash64, the m-mer stream, kmer_view and the
 * generator/mutator -- tests/test_oracle_vs_ref.py compares every function here with the
 * reference's own code executed verbatim (oracle/ref_tier_a.cpp -> oracle/_ref/), and
 * tests/golden/ holds vectors produced that way. For the window-minimum / super-k-mer cut
 * the reference's iterators are non-functional (minimizer_view yields nothing,
 * super_kmer_view does not compile: SURVEY.md F2/F3) and its tests hold no vectors, so
 * that part is "parity unpinned" by the reference: it is a restatement of decisions D1-D9
 *  fed by the pinned per-base stream, with a brute-force and a
 * linear implementation cross-checked against each other.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ nt4 -- */
/* include/constants.hpp:12-21: A,a->0 C,c->1 G,g->2 T,t,U,u->3, everything else 4 */
uint8_t orc_nt4(uint8_t c)
{
    switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': case 'U': case 'u': return 3;
    default: return 4;
    }
}

/* ------------------------------------------------------------ MurmurHash3 -- */
/* bundled/MurmurHash3.cpp:263-341 (Appleby, public domain; published algorithm). */
static inline uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }

static inline uint64_t fmix64(uint64_t h) /* bundled/MurmurHash3.cpp:81-90 */
{
    h ^= h >> 33; h *= 0xff51afd7ed558ccdULL;
    h ^= h >> 33; h *= 0xc4ceb9fe1a85ec53ULL;
    h ^= h >> 33;
    return h;
}

void orc_murmur3_x64_128(const void* key, int len, uint32_t seed, uint64_t out[2])
{
    const uint8_t* p = (const uint8_t*)key;
    const uint64_t c1 = 0x87c37b91114253d5ULL, c2 = 0x4cf5ad432745937fULL;
    uint64_t h1 = seed, h2 = seed;
    int nblocks = len / 16;
    for (int b = 0; b < nblocks; ++b) {
        uint64_t k1, k2;
        memcpy(&k1, p + 16 * b, 8);
        memcpy(&k2, p + 16 * b + 8, 8);
        k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
        h1 = rotl64(h1, 27); h1 += h2; h1 = h1 * 5 + 0x52dce729;
        k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2;
        h2 = rotl64(h2, 31); h2 += h1; h2 = h2 * 5 + 0x38495ab5;
    }
    const uint8_t* tail = p + 16 * nblocks;
    int rem = len & 15;
    uint64_t k1 = 0, k2 = 0;
    for (int i = rem - 1; i >= 8; --i) k2 ^= (uint64_t)tail[i] << (8 * (i - 8));
    if (rem > 8) { k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2; }
    for (int i = (rem > 8 ? 7 : rem - 1); i >= 0; --i) k1 ^= (uint64_t)tail[i] << (8 * i);
    if (rem > 0) { k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1; }
    h1 ^= (uint64_t)len; h2 ^= (uint64_t)len;
    h1 += h2; h2 += h1;
    h1 = fmix64(h1); h2 = fmix64(h2);
    h1 += h2; h2 += h1;
    out[0] = h1; out[1] = h2;
}

/* include/hash.hpp:45-59: key = the 8 object bytes of the value (little endian), seed
 * narrowed to uint32_t, low word of the 128-bit result. */
uint64_t orc_hash64_u64(uint64_t v, uint64_t seed)
{
    uint64_t out[2];
    orc_murmur3_x64_128(&v, 8, (uint32_t)seed, out);
    return out[0];
}
void orc_double_hash64_u64(uint64_t v, uint64_t seed, uint64_t out[2]) { orc_murmur3_x64_128(&v, 8, (uint32_t)seed, out); }
void orc_double_hash64_u128(uint64_t lo, uint64_t hi, uint64_t seed, uint64_t out[2])
{
    uint64_t kk[2] = {lo, hi};
    orc_murmur3_x64_128(kk, 16, (uint32_t)seed, out);
}
/* include/hash.hpp:81-85 */
uint64_t orc_remix(uint64_t z)
{
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}

/* ------------------------------------------------------------ 2-bit words -- */
typedef unsigned __int128 u128;

/* word of the `len` bases starting at s[p] (all assumed valid), first base most
 * significant (include/kmer_view.hpp:194, include/minimizer_view.hpp:236) */
static u128 fwd_word(const char* s, size_t p, unsigned len)
{
    u128 w = 0;
    for (unsigned i = 0; i < len; ++i) w = (w << 2) | orc_nt4((uint8_t)s[p + i]);
    return w;
}
/* reverse complement: complement of the last base most significant
 * (include/kmer_view.hpp:195, include/minimizer_view.hpp:237, from zeroed buffers) */
static u128 rc_word(const char* s, size_t p, unsigned len)
{
    u128 w = 0;
    for (unsigned i = 0; i < len; ++i) w = (w << 2) | (u128)(3 ^ orc_nt4((uint8_t)s[p + len - 1 - i]));
    return w;
}

/* valid_from[i] = number of consecutive valid bases starting at i (capped to the end of the
 * sequence that owns i when offsets are given). Caller frees. */
static uint32_t* run_lengths(const char* s, size_t n, const uint64_t* offsets, size_t n_seq)
{
    uint32_t* r = (uint32_t*)calloc(n + 1, sizeof(uint32_t));
    uint64_t whole[2] = {0, n};
    if (!offsets) { offsets = whole; n_seq = 1; }
    for (size_t q = 0; q < n_seq; ++q) {
        size_t lo = offsets[q], hi = offsets[q + 1];
        for (size_t i = hi; i-- > lo;) {
            uint32_t nxt = (i + 1 == hi) ? 0 : r[i + 1];
            if (orc_nt4((uint8_t)s[i]) < 4) r[i] = nxt < 0xffffffffu ? nxt + 1 : nxt;
            else r[i] = 0;
        }
    }
    return r;
}

/* ---------------------------------------------------------------- kmer_view -- */
/*
 * The record stream of wrapper::kmer_view (include/kmer_view.hpp:155-234), restated over
 * runs of valid bases instead of as an iterator state machine:
 *   - "good run" = maximal run of valid bases of length >= k: [a, b)
 *   - every good run yields its k-mers at positions a .. b-k                 (:181-197)
 *   - between two good runs exactly one record without value, position (b+1)-k, i.e. the
 *     position after the run's last k-mer (K3)                               (:177,187-190,198-200)
 *   - the last good run: if it reaches the end of the sequence its final k-mer is never
 *     yielded (K1, operator!= :57); otherwise all of its k-mers are yielded and nothing
 *     follows (K2 + the K6 extension: input is treated as truncated after the first
 *     invalid byte behind the last good run)
 *   - id = ordinal of the record (K4, :186)
 * bytes >= 0x80 are invalid (K8). Values: canonical ? min(fwd, rc) : fwd. For 64 < 2k the
 * words are 128-bit; canonical then uses the CORRECT reverse complement (the reference's
 * is wrong for k > 32, K7).
 * value_lo/value_hi may be NULL. Returns the number of records (may exceed cap).
 */
size_t orc_kmer_stream(const char* s, size_t n, unsigned k, int canonical,
                       uint64_t* value_lo, uint64_t* value_hi, uint8_t* has_value,
                       uint64_t* position, uint64_t* id, size_t cap)
{
    if (k == 0 || k > 64) return 0;
    size_t cnt = 0;
    /* find the last good run first */
    size_t last_a = 0, last_b = 0; int have = 0;
    {
        size_t run = 0;
        for (size_t i = 0; i <= n; ++i) {
            int v = i < n && orc_nt4((uint8_t)s[i]) < 4;
            if (v) { ++run; continue; }
            if (run >= k) { have = 1; last_a = i - run; last_b = i; }
            run = 0;
        }
    }
    if (!have) return 0;
    size_t run = 0;
    for (size_t i = 0; i <= last_b; ++i) {
        int v = i < last_b && orc_nt4((uint8_t)s[i]) < 4;
        if (i < last_b && v) { ++run; continue; }
        if (run >= k) {
            size_t a = i - run, b = i;
            int is_last = (a == last_a);
            size_t nk = b - a - k + 1;
            if (is_last && b == n) nk -= 1;                       /* K1 */
            for (size_t j = 0; j < nk; ++j) {
                if (cnt < cap) {
                    u128 f = fwd_word(s, a + j, k), val = f;
                    if (canonical) { u128 r = rc_word(s, a + j, k); if (r < f) val = r; }
                    if (value_lo) value_lo[cnt] = (uint64_t)val;
                    if (value_hi) value_hi[cnt] = (uint64_t)(val >> 64);
                    if (has_value) has_value[cnt] = 1;
                    if (position) position[cnt] = a + j;
                    if (id) id[cnt] = cnt;
                }
                ++cnt;
            }
            if (!is_last) {                                       /* K3 */
                if (cnt < cap) {
                    if (value_lo) value_lo[cnt] = 0;
                    if (value_hi) value_hi[cnt] = 0;
                    if (has_value) has_value[cnt] = 0;
                    if (position) position[cnt] = b + 1 - k;
                    if (id) id[cnt] = cnt;
                }
                ++cnt;
            }
        }
        run = 0;
    }
    return cnt;
}

/*
 * The batched form the C-ABI returns (include/biolib_gpu.h blg_kmers): every valid k-mer
 * (all n-k+1 of a clean sequence) as (value, position), positions in the coordinates of the
 * concatenated buffer; sequences delimited by offsets[0..n_seq] (NULL = one sequence).
 */
size_t orc_kmers(const char* s, size_t n, const uint64_t* offsets, size_t n_seq, unsigned k, int canonical,
                 uint64_t* value_lo, uint64_t* value_hi, uint64_t* position, size_t cap)
{
    if (k == 0 || k > 64 || n < k) return 0;
    uint32_t* rl = run_lengths(s, n, offsets, n_seq);
    size_t cnt = 0;
    for (size_t p = 0; p + k <= n; ++p) {
        if (rl[p] < k) continue;
        if (cnt < cap) {
            u128 f = fwd_word(s, p, k), val = f;
            if (canonical) { u128 r = rc_word(s, p, k); if (r < f) val = r; }
            if (value_lo) value_lo[cnt] = (uint64_t)val;
            if (value_hi) value_hi[cnt] = (uint64_t)(val >> 64);
            if (position) position[cnt] = p;
        }
        ++cnt;
    }
    free(rl);
    return cnt;
}

/* ------------------------------------------------------------ m-mer stream -- */
/*
 * D1/D2: for every 0-based start j: valid iff bases [j, j+m) are all valid (and inside one
 * sequence); value = canonical ? min(fwd, rc) : fwd (include/minimizer_view.hpp:236-238 --
 * the strand variable only matters when fwd != rc, so the tie rule D3 never changes the
 * value); hash = hash64(value, seed) (:202,260). m <= 32.
 */
void orc_mmers(const char* s, size_t n, const uint64_t* offsets, size_t n_seq, unsigned m, uint64_t seed, int canonical,
               uint64_t* value, uint64_t* hashv, uint8_t* valid)
{
    uint32_t* rl = run_lengths(s, n, offsets, n_seq);
    for (size_t j = 0; j < n; ++j) {
        int ok = m >= 1 && m <= 32 && rl[j] >= m;
        uint64_t v = 0;
        if (ok) {
            uint64_t f = (uint64_t)fwd_word(s, j, m);
            v = f;
            if (canonical) { uint64_t r = (uint64_t)rc_word(s, j, m); if (r < f) v = r; }
        }
        if (valid) valid[j] = (uint8_t)ok;
        if (value) value[j] = v;
        if (hashv) hashv[j] = ok ? orc_hash64_u64(v, seed) : 0;
    }
    free(rl);
}

/* hash::minimizer_position_extractor (reference include/kmer_view.hpp:266-283): for one k-mer value (k <= 32), the
 * k-m+1 m-mers are visited from the RIGHT (km >>= 2 each step), hashed with seed 0, and a later (more to the left)
 * m-mer replaces the current minimum on '>=' : smallest hash, leftmost on ties. Returned: k - m - minpos, the 0-based
 * offset of that m-mer from the left end of the k-mer. */
void orc_minimizer_positions(const uint64_t* values, size_t n, unsigned k, unsigned m, uint64_t* out)
{
    const uint64_t mask = (2 * m != 64) ? (((uint64_t)1 << (2 * m)) - 1) : ~(uint64_t)0;
    for (size_t q = 0; q < n; ++q) {
        uint64_t km = values[q];
        uint64_t mval = orc_hash64_u64(km & mask, 0);
        unsigned minpos = 0;
        for (unsigned i = 0; i < k - m + 1; ++i) {
            uint64_t val = orc_hash64_u64(km & mask, 0);
            if (mval >= val) { mval = val; minpos = i; }
            km >>= 2;
        }
        out[q] = (uint64_t)(k - m - minpos);
    }
}

/* ------------------------------------------------ minimizers / super-k-mers -- */
/*
 * D4-D8: k-mer at start s valid iff bases [s, s+k) valid and in one
 * sequence; its minimizer = argmin over j in [s, s+w), w = k-m+1, of (hash(j), j) -- smallest
 * hash, leftmost on ties (strict '<' in include/minimizer_view.hpp:283,374). A record starts
 * at the first valid k-mer after a break/sequence start or where the argmin POSITION changes;
 * it spans the consecutive valid k-mers sharing that argmin (intent of
 * include/super_kmer_view.hpp:121-135: mm_pos = minimizer idx - start, size = #k-mers).
 * Output per record: value, hash, mm_start (0-based minimizer start), sk_start (first k-mer),
 * size (#k-mers, <= w). Coordinates are those of the concatenated buffer.
 * mode 0 = brute force O(n*w) (ground truth), mode 1 = monotone deque, linear (the timed CPU arm).
 */
static size_t super_kmers_impl(const char* s, size_t n, const uint64_t* offsets, size_t n_seq,
                               unsigned k, unsigned m, uint64_t seed, int canonical, int mode,
                               uint64_t* value, uint64_t* hashv, uint64_t* mm_start, uint64_t* sk_start,
                               uint8_t* size, size_t cap)
{
    if (m == 0 || m > 32 || k < m || k > 64 || n < k) return 0;
    const unsigned w = k - m + 1;
    uint32_t* rl = run_lengths(s, n, offsets, n_seq);
    size_t cnt = 0;
    /* ring of (hash, value) for the last w m-mers and the deque of candidate positions. 128 slots, not 64: with
     * w = 64 the m-mer at cur_arg + 64 is pushed before the record of cur_arg is closed and must not overwrite it */
    uint64_t* hring = (uint64_t*)malloc(128 * sizeof(uint64_t));
    uint64_t* vring = (uint64_t*)malloc(128 * sizeof(uint64_t));
    size_t dq[128]; unsigned dq_head = 0, dq_len = 0;              /* positions, increasing; hashes strictly increasing front->back */
    uint64_t fw = 0, rc = 0;                                      /* rolling m-mer words */
    const uint64_t mmask = m == 32 ? ~0ULL : ((1ULL << (2 * m)) - 1);
    const unsigned rshift = 2 * (m - 1);
    size_t since_break = 0;                                       /* valid bases since last break */
    int open = 0; size_t cur_arg = 0, cur_start = 0, cur_n = 0;   /* current record */
    size_t seq_i = 0;                                             /* next offsets[] entry to cross; offsets[0]==0, offsets[n_seq]==n */

#define EMIT_RECORD() do { \
        if (cnt < cap) { \
            if (value) value[cnt] = vring[cur_arg & 127]; \
            if (hashv) hashv[cnt] = hring[cur_arg & 127]; \
            if (mm_start) mm_start[cnt] = cur_arg; \
            if (sk_start) sk_start[cnt] = cur_start; \
            if (size) size[cnt] = (uint8_t)cur_n; \
        } ++cnt; open = 0; } while (0)

    if (mode == 1) {
        for (size_t i = 0; i < n; ++i) {
            if (offsets) {                                        /* sequence boundary = break */
                int boundary = 0;
                while (seq_i <= n_seq && offsets[seq_i] == i) { boundary = 1; ++seq_i; }
                if (boundary) { if (open) EMIT_RECORD(); since_break = 0; dq_len = 0; }
            }
            uint8_t c = orc_nt4((uint8_t)s[i]);
            if (c >= 4) {
                if (open) EMIT_RECORD();
                since_break = 0; dq_len = 0;
                continue;
            }
            fw = ((fw << 2) | c) & mmask;
            rc = (rc >> 2) | ((uint64_t)(3 ^ c) << rshift);
            ++since_break;
            if (since_break < m) continue;
            size_t j = i + 1 - m;                                 /* 0-based m-mer start */
            uint64_t v = (canonical && rc < fw) ? rc : fw;
            uint64_t h = orc_hash64_u64(v, seed);
            hring[j & 127] = h; vring[j & 127] = v;
            while (dq_len && hring[dq[(dq_head + dq_len - 1) & 127] & 127] > h) --dq_len;   /* equal stays: leftmost wins */
            dq[(dq_head + dq_len) & 127] = j; ++dq_len;
            if (since_break < k) continue;
            size_t st = i + 1 - k;                                /* k-mer start */
            while (dq[dq_head & 127] < st) { dq_head = (dq_head + 1) & 127; --dq_len; }
            size_t arg = dq[dq_head & 127];
            if (open && arg == cur_arg) { ++cur_n; }
            else {
                if (open) EMIT_RECORD();
                open = 1; cur_arg = arg; cur_start = st; cur_n = 1;
            }
        }
        if (open) EMIT_RECORD();
    } else {
        /* brute force: recompute everything per k-mer from scratch */
        uint64_t* H = (uint64_t*)malloc((n ? n : 1) * sizeof(uint64_t));
        uint64_t* V = (uint64_t*)malloc((n ? n : 1) * sizeof(uint64_t));
        for (size_t j = 0; j < n; ++j) {
            if (rl[j] >= m) {
                uint64_t f = (uint64_t)fwd_word(s, j, m), v = f;
                if (canonical) { uint64_t r = (uint64_t)rc_word(s, j, m); if (r < f) v = r; }
                V[j] = v; H[j] = orc_hash64_u64(v, seed);
            } else { V[j] = 0; H[j] = 0; }
        }
        int prev_valid = 0; size_t prev_arg = 0;
        for (size_t st = 0; st + k <= n; ++st) {
            int ok = rl[st] >= k;
            if (!ok) { if (open) EMIT_RECORD(); prev_valid = 0; continue; }
            size_t arg = st;
            for (size_t j = st + 1; j < st + w; ++j) if (H[j] < H[arg]) arg = j;
            int boundary = 0;
            if (offsets) {                                         /* a k-mer starting a new sequence starts a record */
                size_t lo = 0, hi = n_seq;                          /* is st some offsets[q]? */
                while (lo < hi) { size_t mid = (lo + hi) / 2; if (offsets[mid] < st) lo = mid + 1; else hi = mid; }
                if (lo < n_seq && offsets[lo] == st) boundary = 1;
            }
            if (prev_valid && !boundary && arg == prev_arg) { ++cur_n; }
            else {
                if (open) EMIT_RECORD();
                open = 1; cur_arg = arg; cur_start = st; cur_n = 1;
                hring[arg & 127] = H[arg]; vring[arg & 127] = V[arg];
            }
            prev_valid = 1; prev_arg = arg;
        }
        if (open) EMIT_RECORD();
        free(H); free(V);
    }
#undef EMIT_RECORD
    free(hring); free(vring); free(rl);
    return cnt;
}

size_t orc_super_kmers_brute(const char* s, size_t n, const uint64_t* offsets, size_t n_seq,
                             unsigned k, unsigned m, uint64_t seed, int canonical,
                             uint64_t* value, uint64_t* hashv, uint64_t* mm_start, uint64_t* sk_start, uint8_t* size, size_t cap)
{
    return super_kmers_impl(s, n, offsets, n_seq, k, m, seed, canonical, 0, value, hashv, mm_start, sk_start, size, cap);
}

size_t orc_super_kmers(const char* s, size_t n, const uint64_t* offsets, size_t n_seq,
                       unsigned k, unsigned m, uint64_t seed, int canonical,
                       uint64_t* value, uint64_t* hashv, uint64_t* mm_start, uint64_t* sk_start, uint8_t* size, size_t cap)
{
    return super_kmers_impl(s, n, offsets, n_seq, k, m, seed, canonical, 1, value, hashv, mm_start, sk_start, size, cap);
}

/* --------------------------------------------------- generator / mutator -- */
/* std::mt19937 (the published MT19937 algorithm), as seeded by engine(seed) with the seed
 * narrowed to 32 bits (include/sequence_generator.hpp:17,21). */
typedef struct { uint32_t mt[624]; int idx; } orc_mt;
static void mt_seed(orc_mt* g, uint32_t seed)
{
    g->mt[0] = seed;
    for (int i = 1; i < 624; ++i) g->mt[i] = 1812433253u * (g->mt[i - 1] ^ (g->mt[i - 1] >> 30)) + (uint32_t)i;
    g->idx = 624;
}
static uint32_t mt_next(orc_mt* g)
{
    if (g->idx >= 624) {
        for (int i = 0; i < 624; ++i) {
            uint32_t y = (g->mt[i] & 0x80000000u) | (g->mt[(i + 1) % 624] & 0x7fffffffu);
            g->mt[i] = g->mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        g->idx = 0;
    }
    uint32_t y = g->mt[g->idx++];
    y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
    return y;
}

/* include/sequence_generator.hpp:14-23,34-39: one draw per base through
 * std::uniform_int_distribution<uint8_t>(0,3). With libstdc++ (GCC >= 11) and a 32-bit
 * engine that is Lemire's multiply-shift with range 4 = the top two bits of the draw, never
 * rejecting (2^32 mod 4 == 0). Pinned against the reference run in tests. */
void orc_generate(uint64_t seed, size_t len, char* out)
{
    static const char bases[4] = {'A', 'C', 'G', 'T'};            /* include/constants.hpp:23 */
    orc_mt g; mt_seed(&g, (uint32_t)seed);
    for (size_t i = 0; i < len; ++i) out[i] = bases[mt_next(&g) >> 30];
}

/* std::uniform_real_distribution<double>(0,1) on mt19937 under libstdc++:
 * generate_canonical<double,53> = (d1 + d2 * 2^32) / 2^64 in double, clamped below 1. */
static double mt_unit(orc_mt* g)
{
    double d1 = (double)mt_next(g);
    double d2 = (double)mt_next(g);
    double r = (d1 + d2 * 4294967296.0) / 18446744073709551616.0;
    if (r >= 1.0) r = nextafter(1.0, 0.0);
    return r;
}

/* include/sequence_mutator.hpp:109-169 driven as tests/test_random_sequence_generation.cpp:44-48
 * (absent values dropped). Quirks kept: the last source base is never emitted; a deletion of
 * length L drops one source base (L+1 absent outputs); an insertion repeats the base L extra
 * times. Source must be ACGT only. Returns the output length (may exceed cap). */
size_t orc_mutate(const char* s, size_t n, double mutp, double indelf, double extp, uint64_t seed,
                  char* out, size_t cap, uint64_t counts[3], int64_t* total_indel_len)
{
    static const char bases[4] = {'A', 'C', 'G', 'T'};
    orc_mt g; mt_seed(&g, (uint32_t)seed);
    size_t i = 0, o = 0, indel_len = 0;
    int in_indel = 0, have = 0; char cur = 0;
    uint64_t subs = 0, ins = 0, dels = 0; int64_t tot = 0;
    for (int first = 1;; first = 0) {
        if (first) { if (n == 0) break; }
        else {
            if (i == n) break;                                     /* iterator == cend() */
            if (have) { if (o < cap) out[o] = cur; ++o; }
        }
        /* operator++ */
        if (indel_len == 0) {
            char base = s[i++];
            uint8_t c = orc_nt4((uint8_t)base);
            cur = base; have = 1;
            if (mt_unit(&g) < mutp) {
                if (mt_unit(&g) >= indelf) {
                    c = (uint8_t)((c + (uint8_t)(mt_unit(&g) * 3 + 1)) & 3);
                    ++subs; cur = bases[c];
                } else {
                    indel_len = 1;
                    while (mt_unit(&g) < extp) ++indel_len;
                    in_indel = 1;
                    if (mt_unit(&g) < 0.5) { ++dels; tot -= (int64_t)indel_len; have = 0; }
                    else { ++ins; tot += (int64_t)indel_len; cur = bases[c & 3]; }
                }
            }
        } else {
            if (in_indel) --indel_len;
            if (indel_len == 0) in_indel = 0;
        }
    }
    if (counts) { counts[0] = subs; counts[1] = ins; counts[2] = dels; }
    if (total_indel_len) *total_indel_len = tot;
    return o;
}
