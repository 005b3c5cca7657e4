// jcdf_fock_build_no_alloc — a Fock build for callers that may not allocate (CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
//
// Same contraction as jcdf_fock_build on one dense shard (F = H + 2J - K over B = L^-1 (P|mn), ScreenedDF.jl:227-378 with an
// all-true mask), for SMALL problems only: the half-transformed intermediate W[Q][i][p] — the scratch the handle-based path
// keeps in cudaMalloc'ed shard buffers — lives here in ONE file-scope __device__ array of fixed size, together with the Coulomb
// vector V[Q] and the per-chunk partial Fock matrices.  Limits (include/jcdf.h): N <= 48, Q <= 256, occ <= 16; the array is
// 1 722 368 bytes and every GPU that loads the library holds it until the process exits.
//
//   k_na_half_transform   W[i][p][Q] = sum_n B[Q, p + N n] C[n, i]           lanes over Q (coalesced B columns), one warp per p
//   k_na_coulomb_vector   V[Q] = sum_p sum_i C[p, i] W[i][p][Q]              (= sum_pn B[Q,pn] D[pn], D = C C^T)
//   k_na_tile             Fp[c][p,q] = 2 sum_Q B[Q,pq] V[Q] - sum_Q sum_i W[i][p][Q] W[i][q][Q]   Q in chunk c of 32 rows,
//                         16 x 16 tiles with q-tile >= p-tile, W rows staged in shared memory
//   k_na_sum              F[p,q] = F[q,p] = H + sum_c Fp[c][p,q]             fixed order: bit-reproducible, exactly symmetric
//
// Every word of scratch a call reads was written earlier in the SAME call (W and V for q < Q, Fp for the chunks and tiles in
// use), so nothing has to start at zero and nothing is cleared; the array keeps the last call's contents.  All calls on one
// GPU share the array: calls ordered on one stream are safe, calls that overlap on the GPU (two streams, two host threads,
// two processes' worth of handles in one process) race on it.
#include <mutex>

#include "dmma_common.cuh"
#include "jcdf_macros.h"
#include "../../include/jcdf.h"

namespace jcdf {

constexpr int NA_N = JCDF_NO_ALLOC_MAX_N, NA_Q = JCDF_NO_ALLOC_MAX_Q, NA_OCC = JCDF_NO_ALLOC_MAX_OCC;
constexpr int NA_CHUNK = 32;                     // auxiliary rows per partial Fock matrix
constexpr int NA_CHUNKS = NA_Q / NA_CHUNK;       // 8
constexpr int NA_TILE = 16;
constexpr size_t NA_W_DOUBLES = (size_t)NA_OCC * NA_N * NA_Q;
constexpr size_t NA_V_OFFSET = NA_W_DOUBLES;
constexpr size_t NA_FP_OFFSET = NA_V_OFFSET + NA_Q;
constexpr size_t NA_SCRATCH_DOUBLES = NA_FP_OFFSET + (size_t)NA_CHUNKS * NA_N * NA_N;
static_assert(NA_Q % NA_CHUNK == 0, "whole chunks");

// In host code this name is a host-side stand-in, not a device address: the launcher below never passes it to a kernel, it
// asks the runtime for the device's copy (cudaGetSymbolAddress, once per device) and passes that.
static __device__ __align__(256) double g_no_alloc_scratch[NA_SCRATCH_DOUBLES];

template <int OB>  // occupied orbitals rounded up to 4, 8 or 16: the accumulators stay in registers
__global__ void __launch_bounds__(128) k_na_half_transform(const double* __restrict__ B, int64_t ldb,
                                                            const double* __restrict__ C, int64_t ldc, double* __restrict__ W,
                                                            int N, int Q, int QP, int occ) {
    __shared__ double Cs[NA_N * OB];
    for (int idx = threadIdx.x; idx < N * OB; idx += blockDim.x) {
        const int n = idx / OB, i = idx % OB;
        Cs[idx] = i < occ ? C[n + ldc * i] : 0.0;
    }
    __syncthreads();
    const int q = blockIdx.x * 32 + (threadIdx.x & 31);
    const int p = blockIdx.y * 4 + (threadIdx.x >> 5);
    if (p >= N) return;
    double acc[OB];
#pragma unroll
    for (int i = 0; i < OB; i++) acc[i] = 0.0;
    if (q < Q) {
        const double* col = B + q + ldb * (int64_t)p;
        for (int n = 0; n < N; n++) {
            const double b = __ldg(col + ldb * (int64_t)N * n);
#pragma unroll
            for (int i = 0; i < OB; i++) acc[i] = fma(b, Cs[n * OB + i], acc[i]);
        }
    }
    // rows q in [Q, QP) are written as zeros: the tile kernel reads whole chunks of 32
#pragma unroll
    for (int i = 0; i < OB; i++)
        if (i < occ) W[((int64_t)i * N + p) * QP + q] = acc[i];
}

__global__ void __launch_bounds__(128) k_na_coulomb_vector(const double* __restrict__ C, int64_t ldc,
                                                            const double* __restrict__ W, double* __restrict__ V, int N,
                                                            int QP, int occ) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= QP) return;
    double s = 0.0;
    for (int i = 0; i < occ; i++)
        for (int p = 0; p < N; p++) s = fma(__ldg(C + p + ldc * i), W[((int64_t)i * N + p) * QP + q], s);
    V[q] = s;  // zero for q >= Q (W is zero there)
}

__global__ void __launch_bounds__(NA_TILE * NA_TILE) k_na_tile(const double* __restrict__ B, int64_t ldb,
                                                                const double* __restrict__ W, const double* __restrict__ V,
                                                                double* __restrict__ Fp, int N, int Q, int QP, int occ) {
    __shared__ double Wp[NA_TILE][NA_CHUNK + 1], Wq[NA_TILE][NA_CHUNK + 1], Vs[NA_CHUNK];
    // blockIdx.x -> (tp, tq) with tq >= tp
    const int T = (N + NA_TILE - 1) / NA_TILE;
    int tp = 0, rest = blockIdx.x;
    while (rest >= T - tp) { rest -= T - tp; tp++; }
    const int tq = tp + rest;
    const int chunk = blockIdx.y, q0 = chunk * NA_CHUNK;
    const int tx = threadIdx.x % NA_TILE, ty = threadIdx.x / NA_TILE;
    const int p = tp * NA_TILE + tx, q = tq * NA_TILE + ty;
    if (threadIdx.x < NA_CHUNK) Vs[threadIdx.x] = V[q0 + threadIdx.x];
    double accK = 0.0;
    for (int i = 0; i < occ; i++) {
        __syncthreads();  // previous orbital's tiles consumed (and Vs visible on the first pass)
        for (int idx = threadIdx.x; idx < NA_TILE * NA_CHUNK; idx += NA_TILE * NA_TILE) {
            const int r = idx / NA_CHUNK, k = idx % NA_CHUNK;
            const int pr = tp * NA_TILE + r, qr = tq * NA_TILE + r;
            Wp[r][k] = pr < N ? W[((int64_t)i * N + pr) * QP + q0 + k] : 0.0;
            Wq[r][k] = qr < N ? W[((int64_t)i * N + qr) * QP + q0 + k] : 0.0;
        }
        __syncthreads();
#pragma unroll 8
        for (int k = 0; k < NA_CHUNK; k++) accK = fma(Wp[tx][k], Wq[ty][k], accK);
    }
    if (p >= N || q >= N) return;
    double accJ = 0.0;
    const double* col = B + q0 + ldb * ((int64_t)p + (int64_t)N * q);
    const int nk = min(NA_CHUNK, Q - q0);
    for (int k = 0; k < nk; k++) accJ = fma(__ldg(col + k), Vs[k], accJ);
    Fp[((int64_t)chunk * N + q) * N + p] = 2.0 * accJ - accK;
}

__global__ void __launch_bounds__(256) k_na_sum(const double* __restrict__ Fp, const double* __restrict__ H,
                                                 double* __restrict__ F, int N, int n_chunks) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= N * N) return;
    const int p = idx % N, q = idx / N;
    if (p > q) return;  // the (q, p) element is written by the thread of (p, q): exactly symmetric 2J - K
    double s = 0.0;
    for (int c = 0; c < n_chunks; c++) s += Fp[((int64_t)c * N + q) * N + p];
    F[p + (int64_t)N * q] = s + (H ? H[p + (int64_t)N * q] : 0.0);
    if (p != q) F[q + (int64_t)N * p] = s + (H ? H[q + (int64_t)N * p] : 0.0);
}

// Device address of the scratch array on the CURRENT device, looked up once per device.
static int32_t no_alloc_scratch(double** out) {
    constexpr int MAX_DEV = 64;
    static std::mutex mu;
    static double* ptr[MAX_DEV] = {nullptr};
    int dev = -1;
    JC_CUDA(cudaGetDevice(&dev));
    JC_REQUIRE(dev >= 0 && dev < MAX_DEV, "jcdf_fock_build_no_alloc: device ordinal beyond 63");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {
        double* p = nullptr;
        JC_CUDA(cudaGetSymbolAddress((void**)&p, g_no_alloc_scratch));
        JC_REQUIRE(p != nullptr, "jcdf_fock_build_no_alloc: cudaGetSymbolAddress returned a null address");
        ptr[dev] = p;
    }
    *out = ptr[dev];
    return JCDF_OK;
}

}  // namespace jcdf

using namespace jcdf;

extern "C" int32_t jcdf_fock_build_no_alloc(int64_t N, int64_t Q, int64_t occ, const double* B_dev, int64_t ldb,
                                            const double* C_occ_dev, int64_t ldc, const double* H_dev, double* F_out_dev,
                                            void* stream) {
    JC_REQUIRE(B_dev && C_occ_dev && F_out_dev, "jcdf_fock_build_no_alloc: NULL device pointer (only H_dev may be NULL)");
    if (N < 1 || Q < 1 || occ < 1 || N > NA_N || Q > NA_Q || occ > NA_OCC) {
        set_error("jcdf_fock_build_no_alloc: N = " + std::to_string(N) + ", Q = " + std::to_string(Q) + ", occ = " +
                  std::to_string(occ) + " is outside the fixed scratch array (1 <= N <= " + std::to_string(NA_N) +
                  ", 1 <= Q <= " + std::to_string(NA_Q) + ", 1 <= occ <= " + std::to_string(NA_OCC) +
                  "); use jcdf_create + jcdf_fock_build for larger problems");
        return JCDF_ERR_INVALID;
    }
    JC_REQUIRE(occ <= N, "jcdf_fock_build_no_alloc: occ > N");
    JC_REQUIRE(ldb >= Q && ldc >= N, "jcdf_fock_build_no_alloc: leading dimension smaller than the row count");
    double* scratch = nullptr;
    if (int32_t rc = no_alloc_scratch(&scratch)) return rc;
    double* W = scratch;
    double* V = scratch + NA_V_OFFSET;
    double* Fp = scratch + NA_FP_OFFSET;
    cudaStream_t st = (cudaStream_t)stream;
    const int n = (int)N, q = (int)Q, o = (int)occ;
    const int QP = (q + NA_CHUNK - 1) / NA_CHUNK * NA_CHUNK, n_chunks = QP / NA_CHUNK;
    const dim3 gw(n_chunks, (n + 3) / 4);
    if (o <= 4) k_na_half_transform<4><<<gw, 128, 0, st>>>(B_dev, ldb, C_occ_dev, ldc, W, n, q, QP, o);
    else if (o <= 8) k_na_half_transform<8><<<gw, 128, 0, st>>>(B_dev, ldb, C_occ_dev, ldc, W, n, q, QP, o);
    else k_na_half_transform<16><<<gw, 128, 0, st>>>(B_dev, ldb, C_occ_dev, ldc, W, n, q, QP, o);
    JC_CUDA(cudaGetLastError());
    k_na_coulomb_vector<<<(QP + 127) / 128, 128, 0, st>>>(C_occ_dev, ldc, W, V, n, QP, o);
    JC_CUDA(cudaGetLastError());
    const int T = (n + NA_TILE - 1) / NA_TILE;
    k_na_tile<<<dim3(T * (T + 1) / 2, n_chunks), NA_TILE * NA_TILE, 0, st>>>(B_dev, ldb, W, V, Fp, n, q, QP, o);
    JC_CUDA(cudaGetLastError());
    k_na_sum<<<(n * n + 255) / 256, 256, 0, st>>>(Fp, H_dev, F_out_dev, n, n_chunks);
    JC_CUDA(cudaGetLastError());
    return JCDF_OK;
}
