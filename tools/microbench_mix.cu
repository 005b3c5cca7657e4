// Do DMMA (mma.sync m8n8k4 f64) and DFMA share one FP64 datapath on B200, or can they run concurrently?
// mode 0: every warp DMMA; mode 1: every warp DFMA; mode 2: even warps DMMA, odd warps DFMA (same per-warp work as in
// modes 0/1).  If the pipes are independent, mode 2 finishes in ~max(t0, t1)/2-ish and the summed rate exceeds 37 TF/s.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench_mix microbench_mix.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1);} } while (0)

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(1024) k_mix(double* out, int iters_dmma, int iters_dfma, int mode, double a0, double b0) {
    const int warp = threadIdx.x >> 5;
    const bool do_dmma = (mode == 0) || (mode == 2 && (warp & 1) == 0);
    double s = 0;
    if (do_dmma) {
        double c[8][2];
#pragma unroll
        for (int i = 0; i < 8; i++) { c[i][0] = 0; c[i][1] = 0; }
        double a = a0 + threadIdx.x, b = b0;
        for (int it = 0; it < iters_dmma; it++) {
#pragma unroll
            for (int i = 0; i < 8; i++) dmma884(c[i][0], c[i][1], a, b);
        }
#pragma unroll
        for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1];
    } else {
        double c[8];
#pragma unroll
        for (int i = 0; i < 8; i++) c[i] = i;
        double a = a0 + threadIdx.x * 1e-9, b = b0;
        for (int it = 0; it < iters_dfma; it++) {
#pragma unroll
            for (int i = 0; i < 8; i++) c[i] = fma(c[i], a, b);
        }
#pragma unroll
        for (int i = 0; i < 8; i++) s += c[i];
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    int nsm = prop.multiProcessorCount;
    printf("device %s SMs %d\n", prop.name, nsm);
    double* out; CK(cudaMalloc(&out, sizeof(double) * nsm * 1024));
    const int it_dmma = 20000, it_dfma = 160000;  // equal FMA count per warp: 8*256*20000 == 8*32*160000
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int warps_list[] = {8, 16, 32};
    for (int wi = 0; wi < 3; wi++) {
        int warps = warps_list[wi];
        for (int mode = 0; mode < 3; mode++) {
            float best = 1e30f;
            for (int r = 0; r < 4; r++) {
                cudaEventRecord(e0);
                k_mix<<<nsm, warps * 32>>>(out, it_dmma, it_dfma, mode, 1.0000001, 1e-9);
                cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
                float ms; cudaEventElapsedTime(&ms, e0, e1);
                if (r > 0 && ms < best) best = ms;
            }
            double fl = 2.0 * 8 * 256 * (double)it_dmma * warps * nsm;  // every warp does the same FMA count
            printf("warps/SM %2d mode %d (%s): %8.3f ms  %7.2f TFLOP/s\n", warps, mode,
                   mode == 0 ? "all DMMA" : mode == 1 ? "all DFMA" : "half DMMA + half DFMA", best, fl / best * 1e-9);
        }
    }
    return 0;
}
