// ubench.cu -- instruction-throughput probes for the integer pipes of sm_100a (developer tool, not part of the library).
// Each probe runs an unrolled block of independent operations per thread and reports warp-instructions per clock per SM
// sub-partition at 1, 2 and 4 warps per scheduler, from clock64() deltas. Used to decide which formulation of the
// window-minimum chain / hash / flag packing costs least on the ALU and FMA pipes.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o ubench ubench.cu && ./ubench
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../biolib_b200/csrc/common.cuh"

typedef unsigned int u32;
typedef unsigned long long u64;

constexpr int NACC = 8;      // independent accumulators per thread
constexpr int ITERS = 2000;

struct Res { long long cyc; };

#define PROBE(NAME, OPS_PER_STEP, BODY)                                                                     \
    __global__ void k_##NAME(u32* out, long long* cyc, u32 seed)                                             \
    {                                                                                                        \
        u32 a[NACC], b = seed | 1u, c = seed * 7u + 3u;                                                      \
        for (int i = 0; i < NACC; ++i) a[i] = threadIdx.x * 31u + i * seed;                                  \
        __syncthreads();                                                                                     \
        const long long t0 = clock64();                                                                      \
        for (int it = 0; it < ITERS; ++it) {                                                                 \
            _Pragma("unroll") for (int r = 0; r < 4; ++r) {                                                  \
                _Pragma("unroll") for (int i = 0; i < NACC; ++i) { u32& x = a[i]; BODY }                     \
            }                                                                                                \
        }                                                                                                    \
        const long long t1 = clock64();                                                                      \
        u32 s = 0;                                                                                           \
        for (int i = 0; i < NACC; ++i) s ^= a[i];                                                            \
        out[blockIdx.x * blockDim.x + threadIdx.x] = s ^ b ^ c;                                              \
        if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;                                             \
    }                                                                                                        \
    static const int ops_##NAME = OPS_PER_STEP;

PROBE(lop3, 1, asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(b), "r"(c));)
PROBE(iadd, 1, asm volatile("add.u32 %0, %0, %1;" : "+r"(x) : "r"(b));)
PROBE(iadd3, 1, asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(x) : "r"(b), "r"(c));)
PROBE(imad, 1, asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(b), "r"(c));)
PROBE(imadhi, 1, asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(x) : "r"(b));)
PROBE(shf, 1, asm volatile("shf.r.wrap.b32 %0, %0, %1, 7;" : "+r"(x) : "r"(b));)
PROBE(shr, 1, asm volatile("shr.u32 %0, %0, 1; xor.b32 %0, %0, %1;" : "+r"(x) : "r"(b));)
PROBE(vimnmx, 1, asm volatile("min.u32 %0, %0, %1; add.u32 %0, %0, 5;" : "+r"(x) : "r"(b));)
PROBE(minonly, 1, asm volatile("min.u32 %0, %0, %1;" : "+r"(x) : "r"(c));)
PROBE(setpsel, 2, asm volatile("{.reg .pred p; setp.lt.u32 p, %0, %1; selp.u32 %0, %2, %0, p;}" : "+r"(x) : "r"(b), "r"(c));)
PROBE(prmt, 1, asm volatile("prmt.b32 %0, %0, %1, 0x5140;" : "+r"(x) : "r"(b));)
PROBE(popc, 1, asm volatile("popc.b32 %0, %0;" : "+r"(x));)
PROBE(clz, 1, asm volatile("clz.b32 %0, %0;" : "+r"(x));)
PROBE(brev, 1, asm volatile("brev.b32 %0, %0;" : "+r"(x));)
PROBE(mix_lop_imad, 2, asm volatile("lop3.b32 %0, %0, %1, %2, 0x96; mad.lo.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(b), "r"(c));)
PROBE(mix_lop_imad2, 3,
      asm volatile("lop3.b32 %0, %0, %1, %2, 0x96; mad.lo.u32 %0, %0, %1, %2; mad.lo.u32 %0, %0, %2, %1;" : "+r"(x) : "r"(b), "r"(c));)
PROBE(mix_lop2_imad, 3,
      asm volatile("lop3.b32 %0, %0, %1, %2, 0x96; shf.r.wrap.b32 %0, %0, %1, 9; mad.lo.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(b), "r"(c));)
PROBE(vote, 1, asm volatile("{.reg .pred p; .reg .u32 t; setp.ne.u32 p, %0, 0; vote.sync.ballot.b32 t, p, 0xffffffff; add.u32 %0, %0, t;}" : "+r"(x));)
PROBE(shfl, 1, asm volatile("shfl.sync.down.b32 %0, %0, 1, 0x1f, 0xffffffff;" : "+r"(x));)
PROBE(redux, 1, asm volatile("redux.sync.add.u32 %0, %0, 0xffffffff;" : "+r"(x));)
// tie-flag formulations for one step of a running minimum over packed keys (hash bits above bit 6, index below)
PROBE(chain_packed, 3,
      asm volatile("{.reg .pred p; .reg .u32 t; xor.b32 t, %0, %1; setp.lt.u32 p, t, 64; min.u32 %0, %0, %1; @p or.b32 %2, %2, 1;}"
                   : "+r"(x), "+r"(b), "+r"(c));)
PROBE(chain_packed_c, 3, {
    const u32 k = b + (u32)i;
    c |= (((x ^ k) & 0xFFFFFFC0u) == 0u);
    x = min(x, k);
})
PROBE(chain_free_c, 2, {
    const u32 k = b + (u32)i;
    c |= (x == k);
    x = min(x, k);
})
PROBE(lop3p, 1, asm volatile("{.reg .pred p; .reg .u32 t; xor.b32 t, %0, %1; and.b32 t, t, 0xffffffc0; setp.eq.u32 p, t, 0; @p add.u32 %0, %0, 1;}" : "+r"(x) : "r"(b));)

// 64-bit probes
__global__ void k_mul64(u64* out, long long* cyc, u64 seed)
{
    u64 a[NACC];
    for (int i = 0; i < NACC; ++i) a[i] = threadIdx.x * 31ull + i * seed;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < NACC; ++i) a[i] = a[i] * 0xff51afd7ed558ccdULL;
        }
    }
    const long long t1 = clock64();
    u64 s = 0;
    for (int i = 0; i < NACC; ++i) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
// 64-bit multiply by a constant with the cross terms entering IMAD.WIDE's addend (3 FMA-pipe ops, no separate add)
__device__ __forceinline__ u64 mul64c(u64 a, u32 clo, u32 chi)
{
    const u32 alo = (u32)a, ahi = (u32)(a >> 32);
    u32 t;
    asm("mul.lo.u32 %0, %1, %2;" : "=r"(t) : "r"(ahi), "r"(clo));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(t) : "r"(alo), "r"(chi));
    u64 add, r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(add) : "r"(0u), "r"(t));
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(alo), "r"(clo), "l"(add));
    return r;
}
__global__ void k_mul64w(u64* out, long long* cyc, u64 seed)
{
    u64 a[NACC];
    for (int i = 0; i < NACC; ++i) a[i] = threadIdx.x * 31ull + i * seed;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < NACC; ++i) a[i] = mul64c(a[i], 0xed558ccdu, 0xff51afd7u);
        }
    }
    const long long t1 = clock64();
    u64 s = 0;
    for (int i = 0; i < NACC; ++i) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int VAR>
__global__ void k_hash(u64* out, long long* cyc, u64 seed)
{
    u64 a[NACC];
    for (int i = 0; i < NACC; ++i) a[i] = threadIdx.x * 31ull + i * seed;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < ITERS / 4; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i) a[i] = blg::murmur3_key8(a[i] & 0x3FFFFFFFFFull, 42 ^ 8);
    }
    const long long t1 = clock64();
    u64 s = 0;
    for (int i = 0; i < NACC; ++i) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <typename K, typename T>
static void run(const char* name, K kern, int ops_per_step, T seed, void* out, long long* cyc)
{
    printf("%-18s", name);
    for (int threads : {128, 256, 512, 1024}) {
        long long best = 1LL << 62;
        for (int rep = 0; rep < 3; ++rep) {
            kern<<<148, threads>>>((decltype(seed)*)out, cyc, seed);
            cudaDeviceSynchronize();
            long long c;
            cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
            if (c < best) best = c;
        }
        const double warps_per_smsp = threads / 128.0;
        const double steps = (double)ITERS * 4 * NACC;
        printf("  %2dw/sched: %6.3f inst/clk/smsp (%5.2f clk/step)", (int)warps_per_smsp,
               steps * ops_per_step * warps_per_smsp / best, best / (steps * warps_per_smsp));
    }
    cudaError_t e = cudaGetLastError();
    printf("%s\n", e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main()
{
    void* out;
    long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 8);
    cudaMalloc(&cyc, 8);
#define RUN(NAME) run(#NAME, k_##NAME, ops_##NAME, (u32)12345u, out, cyc)
    RUN(lop3); RUN(iadd); RUN(iadd3); RUN(imad); RUN(imadhi); RUN(shf); RUN(shr); RUN(vimnmx); RUN(minonly); RUN(setpsel);
    RUN(prmt); RUN(popc); RUN(clz); RUN(brev); RUN(mix_lop_imad); RUN(mix_lop_imad2); RUN(mix_lop2_imad);
    RUN(vote); RUN(shfl); RUN(redux); RUN(chain_packed); RUN(chain_packed_c); RUN(chain_free_c); RUN(lop3p);
    run("mul64 (per mul)", k_mul64, 1, (u64)12345u, out, cyc);
    run("mul64w (per mul)", k_mul64w, 1, (u64)12345u, out, cyc);
    // hash: ITERS/4 iterations of NACC hashes: report per-hash clocks by scaling ops so that "clk/step" is per hash
    printf("hash: 'clk/step' below is clocks per 1/16 hash-step; multiply by 16 for clocks per warp-hash per scheduler\n");
    run("murmur3_key8", k_hash<0>, 1, (u64)12345u, out, cyc);
    return 0;
}
