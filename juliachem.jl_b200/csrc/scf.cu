// On-device SCF step around the Fock build: everything scf_cycles_kernel (src/rhf/energy/SCF.jl:340-592)
// does between two calls of df_rhf_fock_build! — DIIS on e = FDS - (FDS)^T (:471-501, DIIS() EnergyHelpers.jl:234-258),
// dynamic damping (:504-508), iteration() (:1072-1125: X'FX, eigen, C = XU, D = 2 C_occ C_occ', E = (D.F + D.H)/2),
// D_rms / dE (:521-525) — plus the one-off orthogonaliser X = S^-1/2 and hcore guess of rhf_kernel (:141-181).
// F, C, D, the DIIS history and every N x N workspace stay resident on the shard-0 device; per iteration the host sees
// ndiis + 3 doubles (the new DIIS Gram row and the three reductions).  The dense N x N algebra is cuBLAS/cuSOLVER (plain
// LAPACK-shaped calls); the hand-written kernels here are the fused reductions and the eigenvector column scaling.
#include <chrono>
#include <cmath>
#include <cstring>
#include <vector>

#include <cublas_v2.h>
#include <cusolverDn.h>

#include "dmma_common.cuh"
#include "jcdf_internal.h"
#include "jcdf_macros.h"

namespace jcdf {

struct ScfState {
    int64_t N = 0;
    int ndiis = 0;
    double E_nuc = 0.0;
    int dev = 0;
    cudaStream_t stream = nullptr;
    cublasHandle_t blas = nullptr;
    cusolverDnHandle_t sol = nullptr;
    // N x N column-major, device
    double *S = nullptr, *X = nullptr, *C = nullptr, *D = nullptr, *Dold = nullptr, *F = nullptr, *Fold = nullptr;
    double *Wa = nullptr, *Wb = nullptr;
    double* evals = nullptr;
    double* work = nullptr;
    int lwork = 0;
    int* dinfo = nullptr;
    double *e_hist = nullptr, *F_hist = nullptr;  // [ndiis][N*N] ring slots (columns of an N^2 x ndiis matrix)
    double* d_small = nullptr;                    // device scratch: [0..ndiis) dots / coefficients, then 3 reductions
    double* d_partial = nullptr;                  // [RED_BLOCKS][3]
    double* h_small = nullptr;                    // pinned mirror of d_small
    std::vector<int> order;                       // history slots, newest first (e_array / F_array of SCF.jl:482-487)
    std::vector<double> gram;                     // [slot][slot] <e_i, e_j>
    int B_dim = 1;
    int iter = 0;  // iterations completed
    double E = 0.0, E_old = 0.0, dE = 1.0, D_rms = 1.0, E_elec = 0.0;
    bool ready = false;
};

constexpr int RED_BLOCKS = 296;
constexpr int RED_THREADS = 256;

// partial[b] = ( sum D*F, sum D*H, sum (D - Dold)^2 ) over a grid-stride slice; fixed order -> deterministic
__global__ void __launch_bounds__(RED_THREADS) k_scf_reduce_partial(const double* __restrict__ D, const double* __restrict__ F,
                                                                     const double* __restrict__ H,
                                                                     const double* __restrict__ Dold, int64_t n,
                                                                     double* __restrict__ partial) {
    double a = 0.0, b = 0.0, c = 0.0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double d = D[i], dd = d - Dold[i];
        a = fma(d, F[i], a);
        b = fma(d, H[i], b);
        c = fma(dd, dd, c);
    }
    __shared__ double red[3][RED_THREADS / 32];
    a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { red[0][warp] = a; red[1][warp] = b; red[2][warp] = c; }
    __syncthreads();
    if (threadIdx.x < 3) {
        double s = 0.0;
        for (int w = 0; w < RED_THREADS / 32; w++) s += red[threadIdx.x][w];
        partial[blockIdx.x * 3 + threadIdx.x] = s;
    }
}
__global__ void k_scf_reduce_final(const double* __restrict__ partial, int nblocks, double* __restrict__ out) {
    if (threadIdx.x < 3) {
        double s = 0.0;
        for (int b = 0; b < nblocks; b++) s += partial[b * 3 + threadIdx.x];
        out[threadIdx.x] = s;
    }
}
// U[:, j] *= (w[j] >= thr ? w[j]^-1/2 : 0)   — Diagonal(S_good)^-0.5 applied to the kept eigenvectors (SCF.jl:153-161)
__global__ void k_scale_columns_invsqrt(double* __restrict__ U, const double* __restrict__ w, int64_t N, double thr) {
    const int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (idx >= N * N) return;
    const int64_t j = idx / N;
    const double ev = w[j];
    U[idx] *= (ev >= thr) ? 1.0 / sqrt(ev) : 0.0;
}

static void scf_free(ScfState* st) {
    if (!st) return;
    cudaSetDevice(st->dev);
    void* ptrs[] = {st->S, st->X, st->C, st->D, st->Dold, st->F, st->Fold, st->Wa, st->Wb, st->evals, st->work, st->dinfo,
                    st->e_hist, st->F_hist, st->d_small, st->d_partial};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    if (st->h_small) cudaFreeHost(st->h_small);
    if (st->blas) cublasDestroy(st->blas);
    if (st->sol) cusolverDnDestroy(st->sol);
    delete st;
}

void scf_destroy(jcdf_handle* h) {
    if (h->scf) scf_free(h->scf);
    h->scf = nullptr;
}

template <typename T>
static int scf_alloc(T** p, size_t count) {
    *p = nullptr;
    JC_CUDA(cudaMalloc(reinterpret_cast<void**>(p), std::max<size_t>(count, 1) * sizeof(T)));
    return JCDF_OK;
}

// iteration() (SCF.jl:1072-1125) on the device: consumes st->F, refreshes C, evals, D (Dold keeps the previous D) and
// leaves ( D.F, D.H, |D - Dold|^2 ) in h_small[ndiis .. ndiis+3) after the stream sync.
static int scf_iteration(jcdf_handle* h, ScfState* st) {
    const int N = (int)st->N;
    const int64_t nn = st->N * st->N;
    const double one = 1.0, zero = 0.0, two = 2.0;
    const double* H = h->shards[0].Hdev;
    JC_CUDA(cudaMemcpyAsync(st->Dold, st->D, sizeof(double) * nn, cudaMemcpyDeviceToDevice, st->stream));
    // workspace_a = X^T F ; workspace_b = workspace_a X      (symm!/gemm!, :1083-1084)
    JC_CUBLAS(cublasDgemm(st->blas, CUBLAS_OP_T, CUBLAS_OP_N, N, N, N, &one, st->X, N, st->F, N, &zero, st->Wa, N));
    JC_CUBLAS(cublasDgemm(st->blas, CUBLAS_OP_N, CUBLAS_OP_N, N, N, N, &one, st->Wa, N, st->X, N, &zero, st->Wb, N));
    // eigen!(Hermitian(workspace_b))  (:1086) — ascending eigenvalues, eigenvectors overwrite Wb
    JC_CUSOLVER(cusolverDnDsyevd(st->sol, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_UPPER, N, st->Wb, N, st->evals, st->work,
                                 st->lwork, st->dinfo));
    // C = X U  (:1095)
    JC_CUBLAS(cublasDgemm(st->blas, CUBLAS_OP_N, CUBLAS_OP_N, N, N, N, &one, st->X, N, st->Wb, N, &zero, st->C, N));
    // D = 2 C_occ C_occ^T  (:1106-1108)
    JC_CUBLAS(cublasDgemm(st->blas, CUBLAS_OP_N, CUBLAS_OP_T, N, N, (int)h->lay.occ, &two, st->C, N, st->C, N, &zero, st->D, N));
    k_scf_reduce_partial<<<RED_BLOCKS, RED_THREADS, 0, st->stream>>>(st->D, st->F, H, st->Dold, nn, st->d_partial);
    k_scf_reduce_final<<<1, 32, 0, st->stream>>>(st->d_partial, RED_BLOCKS, st->d_small + st->ndiis);
    JC_CUDA(cudaGetLastError());
    int info = 0;
    JC_CUDA(cudaMemcpyAsync(&info, st->dinfo, sizeof(int), cudaMemcpyDeviceToHost, st->stream));
    JC_CUDA(cudaMemcpyAsync(st->h_small + st->ndiis, st->d_small + st->ndiis, 3 * sizeof(double), cudaMemcpyDeviceToHost,
                            st->stream));
    JC_CUDA(cudaStreamSynchronize(st->stream));
    if (info != 0) {
        set_error("syevd did not converge in iteration() (info = " + std::to_string(info) + ")");
        return JCDF_ERR_NUMERIC;
    }
    st->E_elec = 0.5 * (st->h_small[st->ndiis] + st->h_small[st->ndiis + 1]);  // (EHF1 + EHF2)/2  (:1113-1115)
    return JCDF_OK;
}

// Solve the (n x n) DIIS system in place (Gaussian elimination, partial pivoting).  Returns false when singular or
// non-finite — the reference's `catch` branch ("Faulty DIIS!", SCF.jl:493-499).
static bool solve_small(std::vector<double>& A, std::vector<double>& b, int n) {
    for (int k = 0; k < n; k++) {
        int piv = k;
        double best = std::fabs(A[k * n + k]);
        for (int r = k + 1; r < n; r++)
            if (std::fabs(A[r * n + k]) > best) { best = std::fabs(A[r * n + k]); piv = r; }
        if (!(best > 0.0) || !std::isfinite(best)) return false;
        if (piv != k) {
            for (int c = 0; c < n; c++) std::swap(A[k * n + c], A[piv * n + c]);
            std::swap(b[k], b[piv]);
        }
        for (int r = k + 1; r < n; r++) {
            const double f = A[r * n + k] / A[k * n + k];
            if (f == 0.0) continue;
            for (int c = k; c < n; c++) A[r * n + c] -= f * A[k * n + c];
            b[r] -= f * b[k];
        }
    }
    for (int k = n - 1; k >= 0; k--) {
        double s = b[k];
        for (int c = k + 1; c < n; c++) s -= A[k * n + c] * b[c];
        b[k] = s / A[k * n + k];
        if (!std::isfinite(b[k])) return false;
    }
    return true;
}

}  // namespace jcdf

using namespace jcdf;

extern "C" {

int32_t jcdf_scf_init(jcdf_handle* h, const double* S, const double* H, double E_nuc, int32_t ndiis, double* E_guess_out) {
    JC_REQUIRE(h && S && H, "jcdf_scf_init: NULL argument");
    JC_REQUIRE(h->lay.mask_set, "jcdf_scf_init: call jcdf_set_mask first");
    JC_REQUIRE(ndiis >= 0 && ndiis <= 64, "jcdf_scf_init: ndiis must be in [0, 64]");
    scf_destroy(h);
    Shard& s0 = h->shards[0];
    const int64_t N = h->lay.N, nn = N * N;
    JC_CUDA(cudaSetDevice(s0.dev));
    ScfState* st = new (std::nothrow) ScfState();
    JC_REQUIRE(st != nullptr, "jcdf_scf_init: out of host memory");
    h->scf = st;
    st->N = N; st->ndiis = ndiis; st->E_nuc = E_nuc; st->dev = s0.dev; st->stream = s0.stream;
    int rc;
    double** mats[] = {&st->S, &st->X, &st->C, &st->D, &st->Dold, &st->F, &st->Fold, &st->Wa, &st->Wb};
    for (double** m : mats)
        if ((rc = scf_alloc(m, (size_t)nn))) return rc;
    if ((rc = scf_alloc(&st->evals, (size_t)N))) return rc;
    if ((rc = scf_alloc(&st->dinfo, 1))) return rc;
    if ((rc = scf_alloc(&st->e_hist, (size_t)nn * std::max(ndiis, 1)))) return rc;
    if ((rc = scf_alloc(&st->F_hist, (size_t)nn * std::max(ndiis, 1)))) return rc;
    if ((rc = scf_alloc(&st->d_small, (size_t)ndiis + 3))) return rc;
    if ((rc = scf_alloc(&st->d_partial, (size_t)RED_BLOCKS * 3))) return rc;
    JC_CUDA(cudaMallocHost(reinterpret_cast<void**>(&st->h_small), sizeof(double) * (ndiis + 3)));
    JC_CUDA(cudaMemsetAsync(st->e_hist, 0, sizeof(double) * nn * std::max(ndiis, 1), st->stream));
    JC_CUDA(cudaMemsetAsync(st->F_hist, 0, sizeof(double) * nn * std::max(ndiis, 1), st->stream));
    JC_CUDA(cudaMemsetAsync(st->D, 0, sizeof(double) * nn, st->stream));
    st->gram.assign((size_t)std::max(ndiis, 1) * std::max(ndiis, 1), 0.0);
    JC_CUBLAS(cublasCreate(&st->blas));
    JC_CUBLAS(cublasSetStream(st->blas, st->stream));
    JC_CUSOLVER(cusolverDnCreate(&st->sol));
    JC_CUSOLVER(cusolverDnSetStream(st->sol, st->stream));
    JC_CUSOLVER(cusolverDnDsyevd_bufferSize(st->sol, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_UPPER, (int)N, st->Wb, (int)N,
                                            st->evals, &st->lwork));
    if ((rc = scf_alloc(&st->work, (size_t)st->lwork))) return rc;
    JC_CUDA(cudaMemcpyAsync(st->S, S, sizeof(double) * nn, cudaMemcpyHostToDevice, st->stream));
    JC_CUDA(cudaMemcpyAsync(s0.Hdev, H, sizeof(double) * nn, cudaMemcpyHostToDevice, st->stream));
    s0.has_H = true;
    // ---- X = U_good diag(s_good^-1/2) U_good^T, eigenvalues below 1e-6 dropped (SCF.jl:141-163) ----
    JC_CUDA(cudaMemcpyAsync(st->Wa, st->S, sizeof(double) * nn, cudaMemcpyDeviceToDevice, st->stream));
    JC_CUSOLVER(cusolverDnDsyevd(st->sol, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_UPPER, (int)N, st->Wa, (int)N, st->evals,
                                 st->work, st->lwork, st->dinfo));
    JC_CUDA(cudaMemcpyAsync(st->Wb, st->Wa, sizeof(double) * nn, cudaMemcpyDeviceToDevice, st->stream));
    k_scale_columns_invsqrt<<<(unsigned)((nn + 255) / 256), 256, 0, st->stream>>>(st->Wb, st->evals, N, 1.0e-6);
    const double one = 1.0, zero = 0.0;
    JC_CUBLAS(cublasDgemm(st->blas, CUBLAS_OP_N, CUBLAS_OP_T, (int)N, (int)N, (int)N, &one, st->Wb, (int)N, st->Wa, (int)N, &zero,
                          st->X, (int)N));
    int info = 0;
    JC_CUDA(cudaMemcpyAsync(&info, st->dinfo, sizeof(int), cudaMemcpyDeviceToHost, st->stream));
    JC_CUDA(cudaStreamSynchronize(st->stream));
    if (info != 0) {
        set_error("syevd of the overlap matrix failed (info = " + std::to_string(info) + ")");
        return JCDF_ERR_NUMERIC;
    }
    // ---- hcore guess: iteration(F = H)  (SCF.jl:178-181); F_old = F (:183) ----
    JC_CUDA(cudaMemcpyAsync(st->F, s0.Hdev, sizeof(double) * nn, cudaMemcpyDeviceToDevice, st->stream));
    if ((rc = scf_iteration(h, st))) return rc;
    JC_CUDA(cudaMemcpyAsync(st->Fold, st->F, sizeof(double) * nn, cudaMemcpyDeviceToDevice, st->stream));
    // block-screened exchange: F is in/out for jcdf_fock_build, and iteration 1 of the reference starts from
    // scf_data.two_electron_fock = zeros (DensityFitting.jl:45), not from the guess matrix
    if (h->lay.xs_any_dropped) JC_CUDA(cudaMemsetAsync(st->F, 0, sizeof(double) * nn, st->stream));
    st->order.clear();
    st->B_dim = 1;
    st->iter = 0;
    st->E = 0.0; st->E_old = 0.0; st->dE = 1.0; st->D_rms = 1.0;
    st->ready = true;
    if (E_guess_out) *E_guess_out = st->E_elec + E_nuc;
    return JCDF_OK;
}

int32_t jcdf_scf_step(jcdf_handle* h, double* result) {
    JC_REQUIRE(h != nullptr, "jcdf_scf_step: NULL handle");
    ScfState* st = h->scf;
    JC_REQUIRE(st && st->ready, "jcdf_scf_step: call jcdf_scf_init first");
    const int N = (int)st->N;
    const int64_t nn = st->N * st->N;
    const double one = 1.0, zero = 0.0, minus_one = -1.0;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto secs = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double>(b - a).count();
    };
    const auto t0 = now();
    const int it = st->iter + 1;
    // ---- F = df_rhf_fock_build!(C)  (SCF.jl:463): C_occ = C[:, 1:occ] and F stay on the device ----
    int rc = fock_build_impl(h, st->C, st->N, nullptr, st->F, nullptr, true);
    if (rc) return rc;
    const auto t1 = now();
    JC_CUDA(cudaSetDevice(st->dev));
    // ---- DIIS (:471-501) ----
    if (st->ndiis > 0) {
        // FDS = F D S ; e = FDS - FDS^T
        JC_CUBLAS(cublasDgemm(st->blas, CUBLAS_OP_N, CUBLAS_OP_N, N, N, N, &one, st->F, N, st->D, N, &zero, st->Wa, N));
        JC_CUBLAS(cublasDgemm(st->blas, CUBLAS_OP_N, CUBLAS_OP_N, N, N, N, &one, st->Wa, N, st->S, N, &zero, st->Wb, N));
        int slot;
        if ((int)st->order.size() < st->ndiis) slot = (int)st->order.size();
        else { slot = st->order.back(); st->order.pop_back(); }
        st->order.insert(st->order.begin(), slot);
        double* e_new = st->e_hist + (int64_t)slot * nn;
        JC_CUBLAS(cublasDgeam(st->blas, CUBLAS_OP_N, CUBLAS_OP_T, N, N, &one, st->Wb, N, &minus_one, st->Wb, N, e_new, N));
        JC_CUDA(cudaMemcpyAsync(st->F_hist + (int64_t)slot * nn, st->F, sizeof(double) * nn, cudaMemcpyDeviceToDevice, st->stream));
        // Gram row of the new error vector against every slot: one gemv over the N^2 x ndiis history matrix
        JC_CUBLAS(cublasDgemv(st->blas, CUBLAS_OP_T, (int)nn, st->ndiis, &one, st->e_hist, (int)nn, e_new, 1, &zero, st->d_small, 1));
        JC_CUDA(cudaMemcpyAsync(st->h_small, st->d_small, sizeof(double) * st->ndiis, cudaMemcpyDeviceToHost, st->stream));
        JC_CUDA(cudaStreamSynchronize(st->stream));
        for (int k = 0; k < st->ndiis; k++) {
            st->gram[(size_t)slot * st->ndiis + k] = st->h_small[k];
            st->gram[(size_t)k * st->ndiis + slot] = st->h_small[k];
        }
        if (it > 1) {
            st->B_dim = std::min(st->B_dim + 1, st->ndiis);
            const int nb = st->B_dim, n = nb + 1;
            std::vector<double> A((size_t)n * n, 0.0), b((size_t)n, 0.0);
            for (int i = 0; i < nb; i++) {
                for (int j = 0; j < nb; j++) A[(size_t)i * n + j] = st->gram[(size_t)st->order[i] * st->ndiis + st->order[j]];
                A[(size_t)i * n + nb] = A[(size_t)nb * n + i] = -1.0;
            }
            b[nb] = -1.0;
            if (solve_small(A, b, n)) {
                for (int k = 0; k < st->ndiis; k++) st->h_small[k] = 0.0;
                for (int i = 0; i < nb; i++) st->h_small[st->order[i]] = b[i];
                JC_CUDA(cudaMemcpyAsync(st->d_small, st->h_small, sizeof(double) * st->ndiis, cudaMemcpyHostToDevice, st->stream));
                // F = sum_i c_i F_i  (EnergyHelpers.jl:252-256): gemv over the N^2 x ndiis Fock history
                JC_CUBLAS(cublasDgemv(st->blas, CUBLAS_OP_N, (int)nn, st->ndiis, &one, st->F_hist, (int)nn, st->d_small, 1, &zero, st->F, 1));
                JC_CUDA(cudaStreamSynchronize(st->stream));  // h_small is reused by iteration()
            } else {
                st->B_dim = 2;  // "Faulty DIIS!" — F keeps this iteration's raw Fock matrix
            }
        }
    }
    // ---- dynamic damping (:504-508): x = 1/log_50(50 dE) when dE >= 1 ----
    const double x = (st->dE >= 1.0) ? 1.0 / (std::log(50.0 * st->dE) / std::log(50.0)) : 1.0;
    if (x != 1.0) {
        const double a = 1.0 - x;
        JC_CUBLAS(cublasDgeam(st->blas, CUBLAS_OP_N, CUBLAS_OP_N, N, N, &a, st->Fold, N, &x, st->F, N, st->Wa, N));
        JC_CUDA(cudaMemcpyAsync(st->F, st->Wa, sizeof(double) * nn, cudaMemcpyDeviceToDevice, st->stream));
    }
    JC_CUDA(cudaMemcpyAsync(st->Fold, st->F, sizeof(double) * nn, cudaMemcpyDeviceToDevice, st->stream));
    const auto t2 = now();
    // ---- new C, D, E (:510-525) ----
    if ((rc = scf_iteration(h, st))) return rc;
    st->D_rms = std::sqrt(st->h_small[st->ndiis + 2]);
    st->E = st->E_elec + st->E_nuc;
    st->dE = st->E - st->E_old;
    st->E_old = st->E;
    st->iter = it;
    const auto t3 = now();
    if (result) {
        result[0] = st->E; result[1] = st->dE; result[2] = st->D_rms; result[3] = st->E_elec;
        result[4] = secs(t0, t1); result[5] = secs(t1, t2); result[6] = secs(t2, t3); result[7] = secs(t0, t3);
    }
    return JCDF_OK;
}

int32_t jcdf_scf_run(jcdf_handle* h, int32_t max_iterations, double dele, double rmsd, double* E_out, int32_t* iterations_out,
                     int32_t* converged_out, double* history) {
    JC_REQUIRE(h != nullptr && max_iterations >= 1, "jcdf_scf_run: bad argument");
    int32_t conv = 0, it = 0;
    double r[8] = {};
    while (it < max_iterations) {
        int rc = jcdf_scf_step(h, r);
        if (rc) return rc;
        if (history) { history[3 * it] = r[0]; history[3 * it + 1] = r[1]; history[3 * it + 2] = r[2]; }
        it++;
        if (std::fabs(r[1]) <= dele && r[2] <= rmsd) { conv = 1; break; }  // SCF.jl:527-530 / :551
    }
    if (E_out) *E_out = r[0];
    if (iterations_out) *iterations_out = it;
    if (converged_out) *converged_out = conv;
    return JCDF_OK;
}

int32_t jcdf_scf_get(jcdf_handle* h, const char* name, double* out) {
    JC_REQUIRE(h && name && out, "jcdf_scf_get: NULL argument");
    ScfState* st = h->scf;
    JC_REQUIRE(st && st->ready, "jcdf_scf_get: call jcdf_scf_init first");
    const std::string n(name);
    const double* src = nullptr;
    size_t count = (size_t)(st->N * st->N);
    if (n == "F") src = st->F;
    else if (n == "D") src = st->D;
    else if (n == "C") src = st->C;
    else if (n == "X") src = st->X;
    else if (n == "S") src = st->S;
    else if (n == "H") src = h->shards[0].Hdev;
    else if (n == "evals") { src = st->evals; count = (size_t)st->N; }
    else { set_error("jcdf_scf_get: unknown matrix '" + n + "' (F, D, C, X, S, H, evals)"); return JCDF_ERR_INVALID; }
    JC_CUDA(cudaSetDevice(st->dev));
    JC_CUDA(cudaMemcpy(out, src, sizeof(double) * count, cudaMemcpyDeviceToHost));
    return JCDF_OK;
}

}  // extern "C"
