// hash64_kernel.cu -- batched hash::hash64::hash<uint64_t> (blg_hash64): reference include/hash.hpp:55-59 ->
// bundled/MurmurHash3.cpp:263-341 for an 8-byte key (murmur3_key8 in common.cuh).
#include "common.cuh"
#include "kernels.h"

namespace blg {

// ---- batched hash::hash64::hash<uint64_t> (include/hash.hpp:55-59) ------------------------------
// Memory-bound (16 B per value against 40 integer instructions): two values per 16-byte access, four accesses in
// flight per thread before the first hash is computed; persistent grid of 148 x 8 CTAs.
__global__ void __launch_bounds__(256) hash64_kernel(const u64* __restrict__ in, u64* __restrict__ out, size_t n, u64 s8)
{
    const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
    const bool vec = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0;
    size_t done = 0;
    if (vec) {
        const ulonglong2* in2 = reinterpret_cast<const ulonglong2*>(in);
        ulonglong2* out2 = reinterpret_cast<ulonglong2*>(out);
        const size_t n2 = n / 2;
        size_t i = tid;
        for (; i + 3 * nth < n2; i += 4 * nth) {
            ulonglong2 v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = __ldg(in2 + i + j * nth);
#pragma unroll
            for (int j = 0; j < 4; ++j) out2[i + j * nth] = make_ulonglong2(murmur3_key8(v[j].x, s8), murmur3_key8(v[j].y, s8));
        }
        for (; i < n2; i += nth) {
            const ulonglong2 v = __ldg(in2 + i);
            out2[i] = make_ulonglong2(murmur3_key8(v.x, s8), murmur3_key8(v.y, s8));
        }
        done = n2 * 2;
    }
    for (size_t i = done + tid; i < n; i += nth) out[i] = murmur3_key8(in[i], s8);
}

cudaError_t hash64_launch(const unsigned long long* in, unsigned long long* out, size_t n, unsigned long long s8,
                          cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    size_t blocks = (n / 2 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    hash64_kernel<<<(unsigned)blocks, 256, 0, stream>>>(in, out, n, s8);
    return cudaGetLastError();
}

}  // namespace blg
