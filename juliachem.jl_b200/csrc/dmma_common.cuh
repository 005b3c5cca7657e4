// Shared device helpers for the sm_100a FP64 tensor-core (DMMA) kernels: mbarrier pipeline, bulk async copies
// (TMA 1-D, SASS UBLKCP), the m8n8k4 FP64 MMA and the tile swizzle used by every HBM-resident operand.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include <cstdio>

// Bounds-checked build (make CHECKED=1 -> libjcdf_checked.so): every global store / bulk-copy source of the pipelined kernels
// is compared with the extent of its buffer; a violation prints the place and traps.  compute-sanitizer is closed on this
// pool, so this build over the smallest case of every kernel family (tools/sanitizer_cases.py) stands in for memcheck.
#ifdef JCDF_CHECKED
#define JC_CHECK(cond, what)                                                                                             \
    do {                                                                                                                 \
        if (!(cond)) {                                                                                                   \
            printf("JCDF_CHECKED violation: %s at %s:%d (block %d, thread %d)\n", what, __FILE__, __LINE__, (int)blockIdx.x, \
                   (int)threadIdx.x);                                                                                    \
            __trap();                                                                                                    \
        }                                                                                                                \
    } while (0)
// Schedule fuzzing of the checked build (no racecheck on this pool): with a non-zero seed every mbarrier wait / arrive, every
// bulk-copy issue and every grid-wide flag operation is preceded, one time in four, by a pseudo-random sleep of up to 4 us that
// depends on the seed, the CTA, the warp and the clock.  The pipelines must then give bit-identical results for every seed;
// a missing ordering shows up as a changed result (tools/sanitizer_cases.py --fuzz N).  One copy of the seed per translation
// unit; JC_DEFINE_FUZZ_SETTER(name) generates the host function that sets it.
static __device__ unsigned int jc_fuzz_seed = 0;
__device__ __forceinline__ void jc_fuzz(unsigned int salt) {
    const unsigned int s = jc_fuzz_seed;
    if (s == 0) return;
    unsigned int h = s ^ (salt * 0x9E3779B1u) ^ (blockIdx.x * 2654435761u) ^ ((threadIdx.x >> 5) * 40503u) ^ (unsigned int)clock();
    h ^= h >> 13;
    h *= 0x5bd1e995u;
    h ^= h >> 15;
    if ((h & 3u) == 0u) __nanosleep((h >> 20) & 0xFFFu);
}
#define JC_FUZZ(salt) jc_fuzz(salt)
#define JC_DEFINE_FUZZ_SETTER(name) \
    void name(unsigned int seed) { cudaMemcpyToSymbol(jc_fuzz_seed, &seed, sizeof(seed)); }
#else
#define JC_CHECK(cond, what) ((void)0)
#define JC_FUZZ(salt) ((void)0)
#define JC_DEFINE_FUZZ_SETTER(name) \
    void name(unsigned int) {}
#endif

namespace jcdf {

constexpr int TILE = 32;                  // edge of a packed B tile (doubles)
constexpr int TILE_ELEMS = TILE * TILE;   // 1024 doubles = 8 KiB, one bulk copy
constexpr int KSLAB = 32;                 // k-depth of one pipeline stage

// Column swizzle applied inside every [32 x W] operand tile stored in HBM (and therefore in shared memory after a
// dense bulk copy): element (row, col) lives at row*W + (col ^ ((row & 3) << 2)).  With it the DMMA fragment loads
// a[m=g][k=t] / b[k=t][n=g] (g = lane/4, t = lane%4) hit 16 distinct 8-byte bank pairs per half warp both when the
// tile is read as [k][m] and when it is read transposed as [m][k].
__host__ __device__ __forceinline__ int swz(int row, int col) { return col ^ ((row & 3) << 2); }

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    JC_FUZZ(1);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    JC_FUZZ(2);
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    JC_FUZZ(3);
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 1-D bulk async copy global -> shared, completion counted in bytes on an mbarrier (TMA unit; SASS UBLKCP).
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    JC_FUZZ(4);
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// D(8x8) += A(8x4) * B(4x8), FP64 tensor core (SASS DMMA.8x8x4).  Fragments per lane (g = lane >> 2, t = lane & 3):
//   a = A[g][t]   b = B[t][g]   c0 = C[g][2t]   c1 = C[g][2t+1]
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

struct PipeState {
    int stage = 0;
    uint32_t phase = 0;
    template <int NSTAGE>
    __device__ __forceinline__ void advance() {
        if (++stage == NSTAGE) { stage = 0; phase ^= 1u; }
    }
};

}  // namespace jcdf
