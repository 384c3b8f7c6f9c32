// shim_bench.cpp -- end-to-end rate of the drop-in a reference user calls: biolib::gpu::super_kmer_view over a std::string,
// every record visited through the iterator (development / BASELINE.md figure; not part of the library).
//   g++ -std=c++17 -O2 -Iinclude tools/shim_bench.cpp -o /tmp/shim_bench -Lbiolib_b200 -lbiolib_gpu -Wl,-rpath,$PWD/biolib_b200
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <string>
#include "biolib_gpu.hpp"
namespace hash { class hash64 {}; }
int main(int argc, char** argv)
{
    const size_t n = argc > 1 ? std::strtoull(argv[1], nullptr, 10) : 100000000ull;
    std::string s(n, 'A');
    blg_synth_generate(42, n, &s[0]);
    for (int want_ids = 1; want_ids >= 0; --want_ids)
        for (int rep = 0; rep < 4; ++rep) {
            const auto t0 = std::chrono::steady_clock::now();
            biolib::gpu::super_kmer_view<uint64_t, uint64_t, hash::hash64> sk(s, 31, 19, true, 42, want_ids != 0);
            auto it = sk.cbegin();                   // the device call happens here
            const double dev_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            uint64_t acc = 0, cnt = 0;
            for (const auto end = sk.cend(); it != end; ++it) { auto const& r = *it; acc += r.minimizer.value + r.minimizer.position + r.size + r.mm_pos; ++cnt; }
            const double call_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            std::printf("%s rep %d: %llu records, checksum %llx, %.2f ms (device call %.2f ms), %.2f Gbases/s (construct + device call + iterate every record)\n",
                        want_ids ? "ids      " : "streaming", rep, (unsigned long long)cnt, (unsigned long long)acc, call_s * 1e3, dev_s * 1e3, n / call_s / 1e9);
        }
    return 0;
}
