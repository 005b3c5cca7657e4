// Gaussian-integral front-end on the GPU: the one-off inputs of the DF Fock build, computed where they
// are consumed instead of on the host with libint + an H2D copy (GPUDF.jl:44-57, :863-864):
//
//   (P|Q)      two-centre metric           replaces calculate_two_center_intgrals   (TwoCenterIntegrals.jl:7-161)
//   (mn|mn)    Schwarz mask + compaction   replaces schwarz_screen_itegrals_df      (SchwarzScreening.jl:9-83)
//   (P|mn)     three-centre integrals      replaces calculate_three_center_integrals (ThreeCenterIntegrals.jl:9-183 and
//              written straight into the   ...Screened.jl) and the C++ engine DFRHFTEIEngine (deps/src/jeri-df-tei.hpp:22-96)
//              tile-packed Q shard
//
// One thread computes one shell block with run-time angular momenta; every launch is uniform in (la, lb, lc).
// FP64 throughout (DFMA pipe); the work is compute-bound and one-off, so the layout goal is simply that the results
// land in the packed HBM layout without a staging copy.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include <cublas_v2.h>

#include "dmma_common.cuh"
#include "integrals.cuh"
#include "jcdf_internal.h"
#include "jcdf_macros.h"

namespace jcdf {

constexpr int INT_THREADS = 64;
constexpr double PI_25_X2 = 34.98683665524972497;  // 2 pi^(5/2)
constexpr int LC_MAX = 6;                         // highest auxiliary angular momentum the scratch arrays admit
constexpr double PRIM_SCREEN = 1e-22;              // |c_a c_b exp(-mu R^2)| below this: the primitive pair is skipped

// doubles of one direction's E array: max over la + lb = Lab of (la+1)(lb+1)(Lab+1)
__host__ __device__ constexpr int edim(int Lab) { return ((Lab + 2) * (Lab + 2) / 4) * (Lab + 1); }

// largest n_A * n_B of a shell pair with la + lb = Lab (balanced split)
__host__ __device__ constexpr int max_nab(int Lab) { return (((Lab + 1) / 2 + 1) * ((Lab + 1) / 2 + 2) / 2) * ((Lab / 2 + 1) * (Lab / 2 + 2) / 2); }

struct IntegralsState {
    struct Basis {
        std::vector<int> l, nprim, prim_off, fn_off;
        std::vector<double> xyz, exps, coefs;
        int nfn = 0;
    } host[2];
    bool have[2] = {false, false};
    // per device copies
    struct Dev {
        int dev = 0;
        BasisDev basis[2];
        std::vector<void*> owned;
        double* boys_table = nullptr;
        int2* d_pairs[4][4] = {};  // primary shell pairs kept by the Schwarz test, by (la, lb) class
        int n_pairs[4][4] = {};
        bool pair_lists_stale = true;
    };
    std::vector<Dev> devs;
    std::vector<uint8_t> shell_mask;  // primary shell pairs kept by the Schwarz test (nshell x nshell), empty = all
    bool in_schwarz = false;          // jcdf_set_mask is being called by jcdf_compute_schwarz_mask (keeps shell_mask)
};

template <typename T>
static int ints_upload(IntegralsState::Dev& d, const std::vector<T>& v, const T** out) {
    T* p = nullptr;
    JC_CUDA(cudaMalloc(reinterpret_cast<void**>(&p), std::max<size_t>(v.size(), 1) * sizeof(T)));
    d.owned.push_back(p);
    JC_CUDA(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    *out = p;
    return JCDF_OK;
}

// jcdf_set_mask hook: a mask that did not come from jcdf_compute_schwarz_mask invalidates the shell-pair list
void integrals_on_set_mask(jcdf_handle* h) {
    IntegralsState* st = h->ints;
    if (!st) return;
    if (!st->in_schwarz) st->shell_mask.clear();
    for (auto& D : st->devs) D.pair_lists_stale = true;
}

void integrals_destroy(jcdf_handle* h) {
    IntegralsState* st = h->ints;
    if (!st) return;
    for (auto& d : st->devs) {
        cudaSetDevice(d.dev);
        for (void* p : d.owned) cudaFree(p);
        for (int la = 0; la < 4; la++)
            for (int lb = 0; lb < 4; lb++)
                if (d.d_pairs[la][lb]) cudaFree(d.d_pairs[la][lb]);
    }
    delete st;
    h->ints = nullptr;
}

// F_n(x) on the host for the interpolation table: ascending series in long double (converges for every x; x <= 40)
static void boys_reference(int nmax, double x, double* out) {
    const long double xl = x, ex = expl(-xl);
    long double term = 1.0L / (2 * nmax + 1), sum = term;
    for (int i = 1; i < 400; i++) {
        term *= 2.0L * xl / (2 * nmax + 2 * i + 1);
        sum += term;
        if (term < 1e-22L * sum) break;
    }
    long double F = ex * sum;
    out[nmax] = (double)F;
    for (int n = nmax; n > 0; n--) {
        F = (2.0L * xl * F + ex) / (2 * n - 1);
        out[n - 1] = (double)F;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// (P|mn): thread = (primary shell pair of one (la, lb) class, auxiliary shell of one lc class)
// ---------------------------------------------------------------------------------------------------------------------
struct ThreeCenterArgs {
    BasisDev prim, aux;
    const int2* pairs;       // primary shell pairs (a >= b in shell order) of this class
    int npairs;
    const int* aux_shells;   // auxiliary shells of this class that overlap the shard
    int naux;
    int la, lb, lc;          // la = l(pairs.x), lb = l(pairs.y)
    const double* boys_table;
    // destination: packed shard
    double* Bp;
    int64_t row_len;
    int64_t q_begin, q_len;
    const int32_t* tile_index;
    const uint8_t* fn_mask;  // N x N or nullptr
    const int32_t* perm;     // caller's function index -> internal (regrouped) index
    int N, nb;
};

// scratch bounds: L_AB <= 6 (f,f), L_TOT <= 10 (f,f,g); ACC doubles of block accumulators per pass over the C components
template <int LAB_MAX, int LTOT_MAX, int ACC>
__global__ void __launch_bounds__(INT_THREADS) k_three_center(ThreeCenterArgs g) {
    const int ip = blockIdx.x * INT_THREADS + threadIdx.x;
    if (ip >= g.npairs) return;
    const int sc = g.aux_shells[blockIdx.y];
    const int2 pr = g.pairs[ip];
    const int sa = pr.x, sb = pr.y;
    const int la = g.la, lb = g.lb, lc = g.lc;
    const int Lab = la + lb, L = Lab + lc;
    const int nA = ncart(la), nB = ncart(lb), nC = ncart(lc), nAB = nA * nB;
    const double Ax = g.prim.xyz[3 * sa], Ay = g.prim.xyz[3 * sa + 1], Az = g.prim.xyz[3 * sa + 2];
    const double Bx = g.prim.xyz[3 * sb], By = g.prim.xyz[3 * sb + 1], Bz = g.prim.xyz[3 * sb + 2];
    const double Cx = g.aux.xyz[3 * sc], Cy = g.aux.xyz[3 * sc + 1], Cz = g.aux.xyz[3 * sc + 2];
    constexpr int EDIM = edim(LAB_MAX);
    double Ex[EDIM], Ey[EDIM], Ez[EDIM];
    double R[(LTOT_MAX + 1) * (LTOT_MAX + 2) * (LTOT_MAX + 3) / 6];
    double G[(LAB_MAX + 1) * (LAB_MAX + 2) * (LAB_MAX + 3) / 6];
    double E1[(LC_MAX + 1) * (LC_MAX + 1)];  // single-Gaussian expansion [i][tau], i, tau <= lc
    double acc[ACC];
    HermIndex ixL, ixab;
    ixL.init(L);
    ixab.init(Lab);
    const int T = Lab + 1, sj = T, si = (lb + 1) * T;
    const int c_chunk = max(1, min(nC, ACC / nAB));
    const double sign = (lc & 1) ? -1.0 : 1.0;  // (-1)^(tau+nu+phi) = (-1)^lc: tau has the parity of c_x, etc.
    const int npa = g.prim.nprim[sa], npb = g.prim.nprim[sb], npc = g.aux.nprim[sc];
    const int oa = g.prim.prim_off[sa], ob = g.prim.prim_off[sb], oc = g.aux.prim_off[sc];
    const int fa0 = g.prim.fn_off[sa], fb0 = g.prim.fn_off[sb], fc0 = g.aux.fn_off[sc];

    for (int c0 = 0; c0 < nC; c0 += c_chunk) {
        const int c1 = min(nC, c0 + c_chunk);
        for (int k = 0; k < (c1 - c0) * nAB; k++) acc[k] = 0.0;
        for (int ka = 0; ka < npa; ka++)
            for (int kb = 0; kb < npb; kb++) {
                const double a = g.prim.exps[oa + ka], b = g.prim.exps[ob + kb];
                const double cc = g.prim.coefs[oa + ka] * g.prim.coefs[ob + kb];
                const double p = a + b;
                const double mu = a * b / p;
                const double r2 = (Ax - Bx) * (Ax - Bx) + (Ay - By) * (Ay - By) + (Az - Bz) * (Az - Bz);
                if (fabs(cc) * exp(-mu * r2) < PRIM_SCREEN) continue;
                hermite_E(la, lb, a, b, Ax, Bx, Ex);
                hermite_E(la, lb, a, b, Ay, By, Ey);
                hermite_E(la, lb, a, b, Az, Bz, Ez);
                const double Px = (a * Ax + b * Bx) / p, Py = (a * Ay + b * By) / p, Pz = (a * Az + b * Bz) / p;
                for (int kc = 0; kc < npc; kc++) {
                    const double q = g.aux.exps[oc + kc];
                    const double alpha = p * q / (p + q);
                    hermite_R(L, alpha, Px - Cx, Py - Cy, Pz - Cz, g.boys_table, ixL, R);
                    const double pref = cc * g.aux.coefs[oc + kc] * sign * PI_25_X2 / (p * q * sqrt(p + q));
                    hermite_E(lc, 0, q, 0.0, 0.0, 0.0, E1);  // E1[i * (lc + 1) + tau]
                    for (int C = c0; C < c1; C++) {
                        int cx, cy, cz;
                        cart_comp(lc, C, cx, cy, cz);
                        // G[tuv] = sum_{tau nu phi} E1[cx,tau] E1[cy,nu] E1[cz,phi] R[t+tau, u+nu, v+phi]
                        for (int t = 0; t <= Lab; t++)
                            for (int u = 0; u <= Lab - t; u++)
                                for (int v = 0; v <= Lab - t - u; v++) {
                                    double s = 0.0;
                                    for (int ta = cx & 1; ta <= cx; ta += 2) {
                                        const double e1 = E1[cx * (lc + 1) + ta];
                                        for (int nu = cy & 1; nu <= cy; nu += 2) {
                                            const double e2 = e1 * E1[cy * (lc + 1) + nu];
                                            for (int ph = cz & 1; ph <= cz; ph += 2)
                                                s = fma(e2 * E1[cz * (lc + 1) + ph], R[ixL(t + ta, u + nu, v + ph)], s);
                                        }
                                    }
                                    G[ixab(t, u, v)] = s;
                                }
                        double* accC = acc + (C - c0) * nAB;
                        for (int A = 0; A < nA; A++) {
                            int ax, ay, az;
                            cart_comp(la, A, ax, ay, az);
                            for (int B = 0; B < nB; B++) {
                                int bx, by, bz;
                                cart_comp(lb, B, bx, by, bz);
                                const double* ex = Ex + ax * si + bx * sj;
                                const double* ey = Ey + ay * si + by * sj;
                                const double* ez = Ez + az * si + bz * sj;
                                double s = 0.0;
                                for (int t = 0; t <= ax + bx; t++) {
                                    double su = 0.0;
                                    for (int u = 0; u <= ay + by; u++) {
                                        double sv = 0.0;
                                        for (int v = 0; v <= az + bz; v++) sv = fma(ez[v], G[ixab(t, u, v)], sv);
                                        su = fma(ey[u], sv, su);
                                    }
                                    s = fma(ex[t], su, s);
                                }
                                accC[A * nB + B] = fma(pref, s, accC[A * nB + B]);
                            }
                        }
                    }
                }
            }
        // ---- scatter the block into the packed shard (component normalisation applied here) ----
        for (int C = c0; C < c1; C++) {
            const int64_t ql = (int64_t)fc0 + C - g.q_begin;
            if (ql < 0 || ql >= g.q_len) continue;
            int cx, cy, cz;
            cart_comp(lc, C, cx, cy, cz);
            const double fC = comp_factor(lc, cx, cy, cz);
            double* row = g.Bp + ql * g.row_len;
            for (int A = 0; A < nA; A++) {
                int ax, ay, az;
                cart_comp(la, A, ax, ay, az);
                const double fA = fC * comp_factor(la, ax, ay, az);
                const int m = fa0 + A;
                for (int B = 0; B < nB; B++) {
                    const int n = fb0 + B;
                    if (n > m) continue;  // same shell: the block is symmetric, (m, n) with m >= n covers it
                    if (g.fn_mask && !g.fn_mask[m + (int64_t)g.N * n]) continue;
                    const int mi = g.perm[m], ni = g.perm[n];
                    const int hi = max(mi, ni), lo = min(mi, ni);
                    const int32_t ti = g.tile_index[(hi >> 5) * g.nb + (lo >> 5)];
                    if (ti < 0) continue;
                    int bx, by, bz;
                    cart_comp(lb, B, bx, by, bz);
                    const double val = fA * comp_factor(lb, bx, by, bz) * acc[(C - c0) * nAB + A * nB + B];
                    const int rm = hi & 31, rn = lo & 31;
                    row[(int64_t)ti * TILE_ELEMS + rm * 32 + swz(rm, rn)] = val;
                    if ((hi >> 5) == (lo >> 5) && hi != lo) row[(int64_t)ti * TILE_ELEMS + rn * 32 + swz(rn, rm)] = val;
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// (P|Q): thread = (auxiliary shell i, auxiliary shell j <= i)
// ---------------------------------------------------------------------------------------------------------------------
struct TwoCenterArgs {
    BasisDev aux;
    const double* boys_table;
    double* J;  // Q x Q column-major, both triangles written
    int Q;
};

__global__ void __launch_bounds__(INT_THREADS) k_two_center(TwoCenterArgs g) {
    const int64_t idx = blockIdx.x * (int64_t)INT_THREADS + threadIdx.x;
    const int ns = g.aux.nshell;
    if (idx >= (int64_t)ns * (ns + 1) / 2) return;
    int si = (int)((sqrt(8.0 * (double)idx + 1.0) - 1.0) * 0.5);
    while ((int64_t)(si + 1) * (si + 2) / 2 <= idx) si++;
    while ((int64_t)si * (si + 1) / 2 > idx) si--;
    const int sj = (int)(idx - (int64_t)si * (si + 1) / 2);
    const int li = g.aux.l[si], lj = g.aux.l[sj], L = li + lj;
    const int ni = ncart(li), nj = ncart(lj);
    const double X = g.aux.xyz[3 * si] - g.aux.xyz[3 * sj], Y = g.aux.xyz[3 * si + 1] - g.aux.xyz[3 * sj + 1],
                 Z = g.aux.xyz[3 * si + 2] - g.aux.xyz[3 * sj + 2];
    double R[(2 * LC_MAX + 1) * (2 * LC_MAX + 2) * (2 * LC_MAX + 3) / 6];
    double Ea[(LC_MAX + 1) * (LC_MAX + 1)], Eb[(LC_MAX + 1) * (LC_MAX + 1)];
    double acc[(LC_MAX + 1) * (LC_MAX + 2) / 2 * ((LC_MAX + 1) * (LC_MAX + 2) / 2)];
    HermIndex ix;
    ix.init(L);
    for (int k = 0; k < ni * nj; k++) acc[k] = 0.0;
    const double sign = (lj & 1) ? -1.0 : 1.0;
    for (int ka = 0; ka < g.aux.nprim[si]; ka++)
        for (int kb = 0; kb < g.aux.nprim[sj]; kb++) {
            const double a = g.aux.exps[g.aux.prim_off[si] + ka], b = g.aux.exps[g.aux.prim_off[sj] + kb];
            const double cc = g.aux.coefs[g.aux.prim_off[si] + ka] * g.aux.coefs[g.aux.prim_off[sj] + kb];
            hermite_R(L, a * b / (a + b), X, Y, Z, g.boys_table, ix, R);
            hermite_E(li, 0, a, 0.0, 0.0, 0.0, Ea);
            hermite_E(lj, 0, b, 0.0, 0.0, 0.0, Eb);
            const double pref = cc * sign * PI_25_X2 / (a * b * sqrt(a + b));
            for (int A = 0; A < ni; A++) {
                int ax, ay, az;
                cart_comp(li, A, ax, ay, az);
                for (int B = 0; B < nj; B++) {
                    int bx, by, bz;
                    cart_comp(lj, B, bx, by, bz);
                    double s = 0.0;
                    for (int t = ax & 1; t <= ax; t += 2)
                        for (int u = ay & 1; u <= ay; u += 2)
                            for (int v = az & 1; v <= az; v += 2) {
                                const double ea = Ea[ax * (li + 1) + t] * Ea[ay * (li + 1) + u] * Ea[az * (li + 1) + v];
                                for (int t2 = bx & 1; t2 <= bx; t2 += 2)
                                    for (int u2 = by & 1; u2 <= by; u2 += 2)
                                        for (int v2 = bz & 1; v2 <= bz; v2 += 2)
                                            s = fma(ea * Eb[bx * (lj + 1) + t2] * Eb[by * (lj + 1) + u2] * Eb[bz * (lj + 1) + v2],
                                                    R[ix(t + t2, u + u2, v + v2)], s);
                            }
                    acc[A * nj + B] = fma(pref, s, acc[A * nj + B]);
                }
            }
        }
    const int fi = g.aux.fn_off[si], fj = g.aux.fn_off[sj];
    for (int A = 0; A < ni; A++) {
        int ax, ay, az;
        cart_comp(li, A, ax, ay, az);
        const double fA = comp_factor(li, ax, ay, az);
        for (int B = 0; B < nj; B++) {
            int bx, by, bz;
            cart_comp(lj, B, bx, by, bz);
            const double v = fA * comp_factor(lj, bx, by, bz) * acc[A * nj + B];
            g.J[(fi + A) + (int64_t)g.Q * (fj + B)] = v;
            g.J[(fj + B) + (int64_t)g.Q * (fi + A)] = v;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Schwarz integrals: thread = primary shell pair (a >= b); block_sum[pair] = sum of the whole (ab|ab) block
// (SchwarzScreening.jl:44-46: a plain sum over all n_mu n_nu x n_mu n_nu entries) and diag[mu + N nu] = (mu nu|mu nu)
// ---------------------------------------------------------------------------------------------------------------------
struct SchwarzArgs {
    BasisDev prim;
    const double* boys_table;
    double* block_sum;  // [nshell (nshell + 1) / 2], index a (a + 1) / 2 + b
    double* diag;       // N x N, both triangles
    int N;
};

template <int LAB_MAX>
__global__ void __launch_bounds__(INT_THREADS) k_schwarz(SchwarzArgs g) {
    const int64_t idx = blockIdx.x * (int64_t)INT_THREADS + threadIdx.x;
    const int ns = g.prim.nshell;
    if (idx >= (int64_t)ns * (ns + 1) / 2) return;
    int sa = (int)((sqrt(8.0 * (double)idx + 1.0) - 1.0) * 0.5);
    while ((int64_t)(sa + 1) * (sa + 2) / 2 <= idx) sa++;
    while ((int64_t)sa * (sa + 1) / 2 > idx) sa--;
    const int sb = (int)(idx - (int64_t)sa * (sa + 1) / 2);
    const int la = g.prim.l[sa], lb = g.prim.l[sb], Lab = la + lb, L = 2 * Lab;
    if (Lab > LAB_MAX) return;  // host checks the bound before launching
    const int nA = ncart(la), nB = ncart(lb), nAB = nA * nB;
    const double Ax = g.prim.xyz[3 * sa], Ay = g.prim.xyz[3 * sa + 1], Az = g.prim.xyz[3 * sa + 2];
    const double Bx = g.prim.xyz[3 * sb], By = g.prim.xyz[3 * sb + 1], Bz = g.prim.xyz[3 * sb + 2];
    constexpr int EDIM = edim(LAB_MAX);
    constexpr int NH = (LAB_MAX + 1) * (LAB_MAX + 2) * (LAB_MAX + 3) / 6;
    double Ex[EDIM], Ey[EDIM], Ez[EDIM], Fx[EDIM], Fy[EDIM], Fz[EDIM];
    double R[(2 * LAB_MAX + 1) * (2 * LAB_MAX + 2) * (2 * LAB_MAX + 3) / 6];
    double Eb[NH], Fb[NH];
    double dacc[max_nab(LAB_MAX)];
    HermIndex ixL, ixab;
    ixL.init(L);
    ixab.init(Lab);
    const int T = Lab + 1, sj = T, si = (lb + 1) * T;
    const int npa = g.prim.nprim[sa], npb = g.prim.nprim[sb];
    const int oa = g.prim.prim_off[sa], ob = g.prim.prim_off[sb];
    const double r2 = (Ax - Bx) * (Ax - Bx) + (Ay - By) * (Ay - By) + (Az - Bz) * (Az - Bz);
    for (int k = 0; k < nAB; k++) dacc[k] = 0.0;
    double total = 0.0;
    const int nh = nherm(Lab);
    for (int k1 = 0; k1 < npa * npb; k1++) {
        const double a = g.prim.exps[oa + k1 / npb], b = g.prim.exps[ob + k1 % npb];
        const double cc = g.prim.coefs[oa + k1 / npb] * g.prim.coefs[ob + k1 % npb];
        const double p = a + b;
        if (fabs(cc) * exp(-(a * b / p) * r2) < PRIM_SCREEN) continue;
        hermite_E(la, lb, a, b, Ax, Bx, Ex);
        hermite_E(la, lb, a, b, Ay, By, Ey);
        hermite_E(la, lb, a, b, Az, Bz, Ez);
        const double Px = (a * Ax + b * Bx) / p, Py = (a * Ay + b * By) / p, Pz = (a * Az + b * Bz) / p;
        // component-summed bra density  Eb[tuv] = sum_AB fA fB E_AB[tuv]
        for (int k = 0; k < nh; k++) Eb[k] = 0.0;
        for (int A = 0; A < nA; A++) {
            int ax, ay, az;
            cart_comp(la, A, ax, ay, az);
            const double fA = comp_factor(la, ax, ay, az);
            for (int B = 0; B < nB; B++) {
                int bx, by, bz;
                cart_comp(lb, B, bx, by, bz);
                const double f = fA * comp_factor(lb, bx, by, bz);
                for (int t = 0; t <= ax + bx; t++)
                    for (int u = 0; u <= ay + by; u++)
                        for (int v = 0; v <= az + bz; v++)
                            Eb[ixab(t, u, v)] += f * Ex[ax * si + bx * sj + t] * Ey[ay * si + by * sj + u] * Ez[az * si + bz * sj + v];
            }
        }
        for (int k2 = 0; k2 < npa * npb; k2++) {
            const double a2 = g.prim.exps[oa + k2 / npb], b2 = g.prim.exps[ob + k2 % npb];
            const double cc2 = g.prim.coefs[oa + k2 / npb] * g.prim.coefs[ob + k2 % npb];
            const double p2 = a2 + b2;
            if (fabs(cc2) * exp(-(a2 * b2 / p2) * r2) < PRIM_SCREEN) continue;
            hermite_E(la, lb, a2, b2, Ax, Bx, Fx);
            hermite_E(la, lb, a2, b2, Ay, By, Fy);
            hermite_E(la, lb, a2, b2, Az, Bz, Fz);
            const double Qx = (a2 * Ax + b2 * Bx) / p2, Qy = (a2 * Ay + b2 * By) / p2, Qz = (a2 * Az + b2 * Bz) / p2;
            hermite_R(L, p * p2 / (p + p2), Px - Qx, Py - Qy, Pz - Qz, g.boys_table, ixL, R);
            const double pref = cc * cc2 * PI_25_X2 / (p * p2 * sqrt(p + p2));
            // ket density with the (-1)^(tau+nu+phi) phase
            for (int k = 0; k < nh; k++) Fb[k] = 0.0;
            for (int A = 0; A < nA; A++) {
                int ax, ay, az;
                cart_comp(la, A, ax, ay, az);
                const double fA = comp_factor(la, ax, ay, az);
                for (int B = 0; B < nB; B++) {
                    int bx, by, bz;
                    cart_comp(lb, B, bx, by, bz);
                    const double f = fA * comp_factor(lb, bx, by, bz);
                    double d = 0.0;  // (AB|AB) of this primitive quartet
                    for (int t2 = 0; t2 <= ax + bx; t2++)
                        for (int u2 = 0; u2 <= ay + by; u2++)
                            for (int v2 = 0; v2 <= az + bz; v2++) {
                                const double ph = ((t2 + u2 + v2) & 1) ? -1.0 : 1.0;
                                const double fk = ph * Fx[ax * si + bx * sj + t2] * Fy[ay * si + by * sj + u2] * Fz[az * si + bz * sj + v2];
                                Fb[ixab(t2, u2, v2)] += f * fk;
                                double s = 0.0;
                                for (int t = 0; t <= ax + bx; t++)
                                    for (int u = 0; u <= ay + by; u++)
                                        for (int v = 0; v <= az + bz; v++)
                                            s = fma(Ex[ax * si + bx * sj + t] * Ey[ay * si + by * sj + u] * Ez[az * si + bz * sj + v],
                                                    R[ixL(t + t2, u + u2, v + v2)], s);
                                d = fma(fk, s, d);
                            }
                    dacc[A * nB + B] = fma(pref * f * f, d, dacc[A * nB + B]);
                }
            }
            double s = 0.0;
            for (int t = 0; t <= Lab; t++)
                for (int u = 0; u <= Lab - t; u++)
                    for (int v = 0; v <= Lab - t - u; v++) {
                        const double eb = Eb[ixab(t, u, v)];
                        if (eb == 0.0) continue;
                        double s2 = 0.0;
                        for (int t2 = 0; t2 <= Lab; t2++)
                            for (int u2 = 0; u2 <= Lab - t2; u2++)
                                for (int v2 = 0; v2 <= Lab - t2 - u2; v2++)
                                    s2 = fma(Fb[ixab(t2, u2, v2)], R[ixL(t + t2, u + u2, v + v2)], s2);
                        s = fma(eb, s2, s);
                    }
            total = fma(pref, s, total);
        }
    }
    g.block_sum[idx] = total;
    const int fa0 = g.prim.fn_off[sa], fb0 = g.prim.fn_off[sb];
    for (int A = 0; A < nA; A++)
        for (int B = 0; B < nB; B++) {
            g.diag[(fa0 + A) + (int64_t)g.N * (fb0 + B)] = dacc[A * nB + B];
            g.diag[(fb0 + B) + (int64_t)g.N * (fa0 + A)] = dacc[A * nB + B];
        }
}

// ---------------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------------
static double dfact_host(int n) {  // n!! with (-1)!! = 1
    double r = 1.0;
    for (; n > 1; n -= 2) r *= n;
    return r;
}

// libint2 Shell::renorm(): primitive norm of the axis-aligned component x^l, then the contraction scaled to unit norm
static void renormalise(int l, int np, const double* exps, const double* raw, double* out) {
    const double pi15 = std::pow(M_PI, 1.5);
    for (int i = 0; i < np; i++)
        out[i] = raw[i] * std::sqrt(std::pow(2.0, l) * std::pow(2.0 * exps[i], l + 1.5) / (pi15 * dfact_host(2 * l - 1)));
    double norm = 0.0;
    for (int i = 0; i < np; i++)
        for (int j = 0; j < np; j++)
            norm += dfact_host(2 * l - 1) * pi15 * out[i] * out[j] / (std::pow(2.0, l) * std::pow(exps[i] + exps[j], l + 1.5));
    const double s = 1.0 / std::sqrt(norm);
    for (int i = 0; i < np; i++) out[i] *= s;
}

static int ensure_devices(jcdf_handle* h, IntegralsState* st) {
    if (!st->devs.empty()) return JCDF_OK;
    std::vector<double> table((size_t)BOYS_GRID * BOYS_ROW);
    for (int g = 0; g < BOYS_GRID; g++) boys_reference(BOYS_ROW - 1, g * BOYS_STEP, table.data() + (size_t)g * BOYS_ROW);
    st->devs.resize(h->shards.size());
    for (size_t d = 0; d < h->shards.size(); d++) {
        IntegralsState::Dev& D = st->devs[d];
        D.dev = h->shards[d].dev;
        JC_CUDA(cudaSetDevice(D.dev));
        const double* t = nullptr;
        int rc = ints_upload(D, table, &t);
        if (rc) return rc;
        D.boys_table = const_cast<double*>(t);
    }
    return JCDF_OK;
}

static int upload_basis(IntegralsState* st, int which) {
    const IntegralsState::Basis& b = st->host[which];
    for (auto& D : st->devs) {
        JC_CUDA(cudaSetDevice(D.dev));
        BasisDev& o = D.basis[which];
        o.nshell = (int)b.l.size();
        o.nfn = b.nfn;
        int rc;
        if ((rc = ints_upload(D, b.l, &o.l))) return rc;
        if ((rc = ints_upload(D, b.nprim, &o.nprim))) return rc;
        if ((rc = ints_upload(D, b.prim_off, &o.prim_off))) return rc;
        if ((rc = ints_upload(D, b.fn_off, &o.fn_off))) return rc;
        if ((rc = ints_upload(D, b.xyz, &o.xyz))) return rc;
        if ((rc = ints_upload(D, b.exps, &o.exps))) return rc;
        if ((rc = ints_upload(D, b.coefs, &o.coefs))) return rc;
    }
    return JCDF_OK;
}

template <int LAB_MAX, int LTOT_MAX, int ACC>
static void launch_three_center_t(const ThreeCenterArgs& a, cudaStream_t st) {
    dim3 grid((unsigned)((a.npairs + INT_THREADS - 1) / INT_THREADS), (unsigned)a.naux);
    k_three_center<LAB_MAX, LTOT_MAX, ACC><<<grid, INT_THREADS, 0, st>>>(a);
}

}  // namespace jcdf

using namespace jcdf;

extern "C" {

int32_t jcdf_set_basis(jcdf_handle* h, int32_t which, int64_t nshell, const int32_t* am, const int32_t* nprim,
                       const double* centers, const double* exponents, const double* coefficients) {
    JC_REQUIRE(h && am && nprim && centers && exponents && coefficients, "jcdf_set_basis: NULL argument");
    JC_REQUIRE(which == 0 || which == 1, "jcdf_set_basis: which must be 0 (primary) or 1 (auxiliary)");
    JC_REQUIRE(nshell > 0 && nshell < (1 << 20), "jcdf_set_basis: bad shell count");
    if (!h->ints) h->ints = new (std::nothrow) IntegralsState();
    JC_REQUIRE(h->ints != nullptr, "jcdf_set_basis: out of host memory");
    IntegralsState* st = h->ints;
    IntegralsState::Basis b;
    int np_total = 0, nf = 0;
    const int lmax = which == 0 ? 3 : 4;
    for (int64_t s = 0; s < nshell; s++) {
        // get_axial_normalization_factor (Globals.jl:30-44) has tables up to am = 5; the kernels here are sized for
        // f primary / g auxiliary functions, the largest the reference's DF inputs use (cc-pVTZ / cc-pVTZ-RIFIT)
        JC_REQUIRE(am[s] >= 0 && am[s] <= lmax, which == 0 ? "jcdf_set_basis: primary angular momentum above f is not supported"
                                                             : "jcdf_set_basis: auxiliary angular momentum above g is not supported");
        JC_REQUIRE(nprim[s] >= 1, "jcdf_set_basis: a shell has no primitives");
        b.l.push_back(am[s]);
        b.nprim.push_back(nprim[s]);
        b.prim_off.push_back(np_total);
        b.fn_off.push_back(nf);
        np_total += nprim[s];
        nf += ncart(am[s]);
        for (int k = 0; k < 3; k++) b.xyz.push_back(centers[3 * s + k]);
    }
    const int64_t expect = which == 0 ? h->lay.N : h->lay.Q;
    JC_REQUIRE(nf == expect, which == 0 ? "jcdf_set_basis: primary shells do not add up to N Cartesian functions"
                                        : "jcdf_set_basis: auxiliary shells do not add up to Q Cartesian functions");
    b.nfn = nf;
    b.exps.assign(exponents, exponents + np_total);
    b.coefs.resize(np_total);
    for (int64_t s = 0; s < nshell; s++) {
        for (int k = 0; k < nprim[s]; k++) JC_REQUIRE(exponents[b.prim_off[s] + k] > 0.0, "jcdf_set_basis: non-positive exponent");
        renormalise(am[s], nprim[s], exponents + b.prim_off[s], coefficients + b.prim_off[s], b.coefs.data() + b.prim_off[s]);
    }
    st->host[which] = b;
    st->have[which] = true;
    for (auto& D : st->devs) D.pair_lists_stale = true;
    int rc = ensure_devices(h, st);
    if (rc) return rc;
    return upload_basis(st, which);
}

int32_t jcdf_compute_two_center(jcdf_handle* h, double* J_out) {
    JC_REQUIRE(h != nullptr, "jcdf_compute_two_center: NULL handle");
    IntegralsState* st = h->ints;
    JC_REQUIRE(st && st->have[1], "jcdf_compute_two_center: call jcdf_set_basis(which = 1) first");
    Shard& s0 = h->shards[0];
    const int64_t Q = h->lay.Q;
    JC_CUDA(cudaSetDevice(s0.dev));
    if (!s0.Metric) JC_CUDA(cudaMalloc(reinterpret_cast<void**>(&s0.Metric), sizeof(double) * Q * Q));
    TwoCenterArgs a;
    a.aux = st->devs[0].basis[1];
    a.boys_table = st->devs[0].boys_table;
    a.J = s0.Metric;
    a.Q = (int)Q;
    const int64_t npairs = (int64_t)a.aux.nshell * (a.aux.nshell + 1) / 2;
    k_two_center<<<(unsigned)((npairs + INT_THREADS - 1) / INT_THREADS), INT_THREADS, 0, s0.stream>>>(a);
    JC_CUDA(cudaGetLastError());
    JC_CUDA(cudaStreamSynchronize(s0.stream));
    h->metric_on_device = true;
    h->host_metric.clear();
    std::vector<double> diagJ((size_t)Q);
    JC_CUDA(cudaMemcpy2D(diagJ.data(), sizeof(double), s0.Metric, sizeof(double) * (Q + 1), sizeof(double), (size_t)Q,
                         cudaMemcpyDeviceToHost));
    h->max_PP = 0.0;
    for (double v : diagJ) h->max_PP = std::max(h->max_PP, std::fabs(v));
    if (J_out) JC_CUDA(cudaMemcpy(J_out, s0.Metric, sizeof(double) * Q * Q, cudaMemcpyDeviceToHost));
    return JCDF_OK;
}

int32_t jcdf_compute_schwarz_mask(jcdf_handle* h, double sigma, uint8_t* fn_mask_out, int64_t* pairs_kept_out) {
    JC_REQUIRE(h != nullptr, "jcdf_compute_schwarz_mask: NULL handle");
    IntegralsState* st = h->ints;
    JC_REQUIRE(st && st->have[0], "jcdf_compute_schwarz_mask: call jcdf_set_basis(which = 0) first");
    const int64_t N = h->lay.N, Q = h->lay.Q;
    const int ns = (int)st->host[0].l.size();
    st->shell_mask.clear();
    for (auto& D : st->devs) D.pair_lists_stale = true;
    if (!(sigma > 0.0)) {  // unscreened: setup_unscreened_screening_matricies (SchwarzScreening.jl:97-111)
        if (fn_mask_out) memset(fn_mask_out, 1, (size_t)N * N);
        if (pairs_kept_out) *pairs_kept_out = N * N;
        st->in_schwarz = true;
        const int rc0 = jcdf_set_mask(h, nullptr, nullptr);
        st->in_schwarz = false;
        return rc0;
    }
    JC_REQUIRE(h->max_PP > 0.0, "jcdf_compute_schwarz_mask: the two-centre metric is needed first (max_P (P|P))");
    Shard& s0 = h->shards[0];
    JC_CUDA(cudaSetDevice(s0.dev));
    const double maxPP = h->max_PP;  // get_max_P_P (SchwarzScreening.jl:86-95)
    const double threshold = sigma * sigma / maxPP;  // :22
    const int64_t npairs = (int64_t)ns * (ns + 1) / 2;
    double *d_sum = nullptr, *d_diag = nullptr;
    JC_CUDA(cudaMalloc(reinterpret_cast<void**>(&d_sum), sizeof(double) * npairs));
    JC_CUDA(cudaMalloc(reinterpret_cast<void**>(&d_diag), sizeof(double) * N * N));
    SchwarzArgs a;
    a.prim = st->devs[0].basis[0];
    a.boys_table = st->devs[0].boys_table;
    a.block_sum = d_sum;
    a.diag = d_diag;
    a.N = (int)N;
    int lmax = 0;
    for (int l : st->host[0].l) lmax = std::max(lmax, l);
    const unsigned grid = (unsigned)((npairs + INT_THREADS - 1) / INT_THREADS);
    if (2 * lmax <= 4) k_schwarz<4><<<grid, INT_THREADS, 0, s0.stream>>>(a);
    else k_schwarz<6><<<grid, INT_THREADS, 0, s0.stream>>>(a);
    JC_CUDA(cudaGetLastError());
    std::vector<double> sum((size_t)npairs), diag((size_t)N * N);
    JC_CUDA(cudaMemcpyAsync(sum.data(), d_sum, sizeof(double) * npairs, cudaMemcpyDeviceToHost, s0.stream));
    JC_CUDA(cudaMemcpyAsync(diag.data(), d_diag, sizeof(double) * N * N, cudaMemcpyDeviceToHost, s0.stream));
    JC_CUDA(cudaStreamSynchronize(s0.stream));
    cudaFree(d_sum);
    cudaFree(d_diag);
    // shell-pair test on the block sum, function-pair test on the diagonal (SchwarzScreening.jl:46-70)
    const IntegralsState::Basis& b = st->host[0];
    std::vector<uint8_t> mask((size_t)N * N, 0);
    st->shell_mask.assign((size_t)ns * ns, 0);
    for (int i = 0; i < ns; i++)
        for (int j = 0; j <= i; j++) {
            const bool keep = !(std::fabs(sum[(size_t)i * (i + 1) / 2 + j]) < threshold);
            st->shell_mask[(size_t)i * ns + j] = st->shell_mask[(size_t)j * ns + i] = keep;
            if (!keep) continue;
            for (int m = b.fn_off[i]; m < b.fn_off[i] + ncart(b.l[i]); m++)
                for (int n = b.fn_off[j]; n < b.fn_off[j] + ncart(b.l[j]); n++) {
                    const bool k = !(std::fabs(diag[(size_t)m + N * n]) < threshold);
                    mask[(size_t)m + N * n] = mask[(size_t)n + N * m] = k;
                }
        }
    // pair compaction: outer loop p (column), inner loop q (row) over kept pairs (:72-81)
    std::vector<int64_t> map((size_t)N * N, 0);
    int64_t idx = 1;
    for (int64_t p = 0; p < N; p++)
        for (int64_t q = 0; q < N; q++)
            if (mask[(size_t)q + N * p]) map[(size_t)q + N * p] = idx++;
    if (fn_mask_out) memcpy(fn_mask_out, mask.data(), (size_t)N * N);
    if (pairs_kept_out) *pairs_kept_out = idx - 1;
    st->in_schwarz = true;
    const int rc1 = jcdf_set_mask(h, mask.data(), map.data());
    st->in_schwarz = false;
    return rc1;
}

// (P|mn) for the auxiliary functions [q_begin, q_begin + q_len) into `dst` ([q_len][row_len], packed layout, zeroed
// here) on device slot d of the handle; all launches go to the shard's stream, which is synchronised before returning.
static int three_center_rows(jcdf_handle* h, size_t d, int64_t q_begin, int64_t q_len, double* dst) {
    IntegralsState* st = h->ints;
    const Layout& L = h->lay;
    const IntegralsState::Basis& pb = st->host[0];
    const IntegralsState::Basis& ab = st->host[1];
    const int ns = (int)pb.l.size();
    Shard& s = h->shards[d];
    IntegralsState::Dev& D = st->devs[d];
    JC_CUDA(cudaSetDevice(s.dev));
    JC_CUDA(cudaMemsetAsync(dst, 0, sizeof(double) * (size_t)q_len * (size_t)L.row_len, s.stream));
    if (q_len == 0) return JCDF_OK;
    // primary shell pairs by (la, lb) class; pairs the Schwarz test dropped as a whole are never computed
    if (D.pair_lists_stale) {
        for (int la = 0; la < 4; la++)
            for (int lb = 0; lb < 4; lb++) {
                if (D.d_pairs[la][lb]) cudaFree(D.d_pairs[la][lb]);
                D.d_pairs[la][lb] = nullptr;
                D.n_pairs[la][lb] = 0;
            }
        std::vector<int2> cls[4][4];
        for (int i = 0; i < ns; i++)
            for (int j = 0; j <= i; j++) {
                if (!st->shell_mask.empty() && !st->shell_mask[(size_t)i * ns + j]) continue;
                cls[pb.l[i]][pb.l[j]].push_back(make_int2(i, j));
            }
        for (int la = 0; la < 4; la++)
            for (int lb = 0; lb < 4; lb++) {
                const std::vector<int2>& pr = cls[la][lb];
                if (pr.empty()) continue;
                JC_CUDA(cudaMalloc(reinterpret_cast<void**>(&D.d_pairs[la][lb]), sizeof(int2) * pr.size()));
                JC_CUDA(cudaMemcpy(D.d_pairs[la][lb], pr.data(), sizeof(int2) * pr.size(), cudaMemcpyHostToDevice));
                D.n_pairs[la][lb] = (int)pr.size();
            }
        D.pair_lists_stale = false;
    }
    std::vector<int> aux_by_l[5];
    for (size_t c = 0; c < ab.l.size(); c++) {
        const int64_t f0 = ab.fn_off[c], f1 = f0 + ncart(ab.l[c]);
        if (f1 > q_begin && f0 < q_begin + q_len) aux_by_l[ab.l[c]].push_back((int)c);
    }
    std::vector<void*> tmp;
    for (int lc = 0; lc < 5; lc++) {
        const std::vector<int>& ax = aux_by_l[lc];
        if (ax.empty()) continue;
        int* d_aux = nullptr;
        JC_CUDA(cudaMalloc(reinterpret_cast<void**>(&d_aux), sizeof(int) * ax.size()));
        tmp.push_back(d_aux);
        JC_CUDA(cudaMemcpyAsync(d_aux, ax.data(), sizeof(int) * ax.size(), cudaMemcpyHostToDevice, s.stream));
        for (int la = 0; la < 4; la++)
            for (int lb = 0; lb < 4; lb++) {
                if (D.n_pairs[la][lb] == 0) continue;
                ThreeCenterArgs a;
                a.prim = D.basis[0];
                a.aux = D.basis[1];
                a.pairs = D.d_pairs[la][lb];
                a.npairs = D.n_pairs[la][lb];
                a.aux_shells = d_aux;
                a.naux = (int)ax.size();
                a.la = la; a.lb = lb; a.lc = lc;
                a.boys_table = D.boys_table;
                a.Bp = dst;
                a.row_len = L.row_len;
                a.q_begin = q_begin;
                a.q_len = q_len;
                a.tile_index = s.d_tile_index;
                a.fn_mask = s.d_fn_mask;
                a.perm = s.d_perm;
                a.N = (int)L.N;
                a.nb = L.nb;
                const int Lab = la + lb, Lt = Lab + lc;
                if (Lab <= 2 && Lt <= 6) launch_three_center_t<2, 6, 128>(a, s.stream);
                else if (Lab <= 4 && Lt <= 8) launch_three_center_t<4, 8, 384>(a, s.stream);
                else launch_three_center_t<6, 10, 400>(a, s.stream);
                JC_CUDA(cudaGetLastError());
            }
    }
    JC_CUDA(cudaStreamSynchronize(s.stream));
    for (void* p : tmp) cudaFree(p);
    return JCDF_OK;
}

int32_t jcdf_compute_three_center(jcdf_handle* h) {
    JC_REQUIRE(h != nullptr, "jcdf_compute_three_center: NULL handle");
    IntegralsState* st = h->ints;
    JC_REQUIRE(st && st->have[0] && st->have[1], "jcdf_compute_three_center: call jcdf_set_basis for both basis sets first");
    JC_REQUIRE(h->lay.mask_set, "jcdf_compute_three_center: call jcdf_set_mask or jcdf_compute_schwarz_mask first");
    for (size_t d = 0; d < h->shards.size(); d++) {
        Shard& s = h->shards[d];
        int rc = three_center_rows(h, d, s.q_begin, s.q_len, s.Bp);
        if (rc) return rc;
        s.t_uploaded = true;
        s.b_ready = false;
        s.bc_valid = false;
    }
    return JCDF_OK;
}

// B = L^-1 (P|mn) per device without any exchange between devices or ranks: every device factors its
// own copy of the (cheap) metric and RECOMPUTES the raw integrals of the auxiliary rows h <= its shard in chunks,
//     B_g = sum_h Linv[Q_g, Q_h] T_h ,
// instead of staging T blocks through host memory (calculate_B_GPU!, GPUDF.jl:918-997; ScreenedDF.jl:134-190).  The
// three-centre integrals of all of w50 cost 0.25 s on one B200 — far less than moving them — so this also covers
// one-process-per-GPU runs, where jcdf_factor_and_form_B cannot see the other ranks' T.
int32_t jcdf_compute_B(jcdf_handle* h) {
    JC_REQUIRE(h != nullptr, "jcdf_compute_B: NULL handle");
    IntegralsState* st = h->ints;
    JC_REQUIRE(st && st->have[0] && st->have[1], "jcdf_compute_B: call jcdf_set_basis for both basis sets first");
    const Layout& L = h->lay;
    JC_REQUIRE(L.mask_set, "jcdf_compute_B: call jcdf_set_mask or jcdf_compute_schwarz_mask first");
    const int64_t Q = L.Q;
    const double one = 1.0, zero = 0.0;
    for (size_t d = 0; d < h->shards.size(); d++) {
        Shard& s = h->shards[d];
        JC_CUDA(cudaSetDevice(s.dev));
        if (s.q_len == 0) { s.b_ready = true; s.bc_valid = false; continue; }
        // ---- metric, potrf('L'), Linv on this device ----
        if (!s.Metric) JC_CUDA(cudaMalloc(reinterpret_cast<void**>(&s.Metric), sizeof(double) * Q * Q));
        if (!s.Linv) JC_CUDA(cudaMalloc(reinterpret_cast<void**>(&s.Linv), sizeof(double) * Q * Q));
        TwoCenterArgs ta;
        ta.aux = st->devs[d].basis[1];
        ta.boys_table = st->devs[d].boys_table;
        ta.J = s.Metric;
        ta.Q = (int)Q;
        const int64_t npairs = (int64_t)ta.aux.nshell * (ta.aux.nshell + 1) / 2;
        k_two_center<<<(unsigned)((npairs + INT_THREADS - 1) / INT_THREADS), INT_THREADS, 0, s.stream>>>(ta);
        JC_CUDA(cudaGetLastError());
        if (d == 0) h->metric_on_device = false;  // consumed below
        int rc = factor_metric_on_device(s, Q);
        if (rc) return rc;
        // ---- B_g += Linv[Q_g, chunk] T_chunk over the rows h < q_begin + q_len ----
        const int64_t q_end = s.q_begin + s.q_len;
        int64_t R = std::max<int64_t>(32, std::min<int64_t>(q_end, ((int64_t)3 << 27) / std::max<int64_t>(L.row_len, 1)));  // <= 3 GiB
        struct Scratch {  // released on every exit path
            double* T = nullptr;
            cublasHandle_t blas = nullptr;
            ~Scratch() { if (blas) cublasDestroy(blas); if (T) cudaFree(T); }
        } sc;
        JC_CUDA(cudaMalloc(reinterpret_cast<void**>(&sc.T), sizeof(double) * (size_t)R * (size_t)L.row_len));
        double* const T = sc.T;
        JC_CUBLAS(cublasCreate(&sc.blas));
        cublasHandle_t blas = sc.blas;
        JC_CUBLAS(cublasSetStream(blas, s.stream));
        bool first = true;
        for (int64_t h0 = 0; h0 < q_end; h0 += R) {
            const int64_t rows = std::min(R, q_end - h0);
            if ((rc = three_center_rows(h, d, h0, rows, T))) return rc;
            // packed rows [q][row_len] viewed column-major are (row_len x q): Bp^T += T^T Linv[Q_g, chunk]^T, in column
            // windows so that every GEMM dimension fits cuBLAS' 32-bit interface
            // Linv is lower triangular (exact zeros above the diagonal: trsm against the identity), so the chunk only
            // reaches the shard rows Q >= h0 — about half of the rectangle on a single device
            const int64_t q_lo = std::max<int64_t>(0, h0 - s.q_begin);
            for (int64_t c0 = 0; c0 < L.row_len && q_lo < s.q_len; c0 += (int64_t)1 << 30) {
                const int64_t w = std::min<int64_t>((int64_t)1 << 30, L.row_len - c0);
                JC_CUBLAS(cublasDgemm(blas, CUBLAS_OP_N, CUBLAS_OP_T, (int)w, (int)(s.q_len - q_lo), (int)rows, &one, T + c0,
                                      (int)L.row_len, s.Linv + s.q_begin + q_lo + Q * h0, (int)Q, first ? &zero : &one,
                                      s.Bp + c0 + q_lo * L.row_len, (int)L.row_len));
            }
            first = false;
        }
        JC_CUDA(cudaStreamSynchronize(s.stream));
        s.t_uploaded = false;
        s.b_ready = true;
        s.bc_valid = false;
    }
    return JCDF_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// One-electron integrals S, T, V (what rhf_kernel feeds the SCF with, src/rhf/energy/SCF.jl:100-135; libint one-body
// engines behind deps/src/jeri-oei.hpp): thread = primary shell pair (a >= b).  Needed by the resident SCF loop so a
// whole DF-RHF run can start from (geometry, basis sets) without any host integral code.
// ---------------------------------------------------------------------------------------------------------------------
namespace jcdf {

struct OneElectronArgs {
    BasisDev prim;
    const double* boys_table;
    const double* charges;  // [natoms]
    const double* nuclei;   // [3 * natoms]
    int natoms;
    double *S, *T, *V;      // N x N column-major, both triangles written
    int N;
};

template <int LAB_MAX>
__global__ void __launch_bounds__(INT_THREADS) k_one_electron(OneElectronArgs g) {
    const int64_t idx = blockIdx.x * (int64_t)INT_THREADS + threadIdx.x;
    const int ns = g.prim.nshell;
    if (idx >= (int64_t)ns * (ns + 1) / 2) return;
    int sa = (int)((sqrt(8.0 * (double)idx + 1.0) - 1.0) * 0.5);
    while ((int64_t)(sa + 1) * (sa + 2) / 2 <= idx) sa++;
    while ((int64_t)sa * (sa + 1) / 2 > idx) sa--;
    const int sb = (int)(idx - (int64_t)sa * (sa + 1) / 2);
    const int la = g.prim.l[sa], lb = g.prim.l[sb], Lab = la + lb;
    if (Lab > LAB_MAX) return;
    const int nA = ncart(la), nB = ncart(lb), nAB = nA * nB;
    const double Ax = g.prim.xyz[3 * sa], Ay = g.prim.xyz[3 * sa + 1], Az = g.prim.xyz[3 * sa + 2];
    const double Bx = g.prim.xyz[3 * sb], By = g.prim.xyz[3 * sb + 1], Bz = g.prim.xyz[3 * sb + 2];
    constexpr int EDIM = edim(LAB_MAX + 2);
    double Ex[EDIM], Ey[EDIM], Ez[EDIM];
    double R[(LAB_MAX + 1) * (LAB_MAX + 2) * (LAB_MAX + 3) / 6], Rz[(LAB_MAX + 1) * (LAB_MAX + 2) * (LAB_MAX + 3) / 6];
    double accS[max_nab(LAB_MAX)], accT[max_nab(LAB_MAX)], accV[max_nab(LAB_MAX)];
    HermIndex ix;
    ix.init(Lab);
    const int lb2 = lb + 2, T = la + lb2 + 1, sj = T, si = (lb2 + 1) * T;
    const int nh = nherm(Lab);
    for (int k = 0; k < nAB; k++) accS[k] = accT[k] = accV[k] = 0.0;
    const int npa = g.prim.nprim[sa], npb = g.prim.nprim[sb];
    const int oa = g.prim.prim_off[sa], ob = g.prim.prim_off[sb];
    const double r2 = (Ax - Bx) * (Ax - Bx) + (Ay - By) * (Ay - By) + (Az - Bz) * (Az - Bz);
    for (int ka = 0; ka < npa; ka++)
        for (int kb = 0; kb < npb; kb++) {
            const double a = g.prim.exps[oa + ka], b = g.prim.exps[ob + kb];
            const double cc = g.prim.coefs[oa + ka] * g.prim.coefs[ob + kb];
            const double p = a + b;
            if (fabs(cc) * exp(-(a * b / p) * r2) < PRIM_SCREEN) continue;
            hermite_E(la, lb2, a, b, Ax, Bx, Ex);
            hermite_E(la, lb2, a, b, Ay, By, Ey);
            hermite_E(la, lb2, a, b, Az, Bz, Ez);
            const double Px = (a * Ax + b * Bx) / p, Py = (a * Ay + b * By) / p, Pz = (a * Az + b * Bz) / p;
            for (int k = 0; k < nh; k++) Rz[k] = 0.0;
            for (int c = 0; c < g.natoms; c++) {
                hermite_R(Lab, p, Px - g.nuclei[3 * c], Py - g.nuclei[3 * c + 1], Pz - g.nuclei[3 * c + 2], g.boys_table, ix, R);
                const double z = g.charges[c];
                for (int k = 0; k < nh; k++) Rz[k] = fma(z, R[k], Rz[k]);
            }
            const double rt = sqrt(3.14159265358979323846 / p), rt3 = rt * rt * rt;
            const double vpref = -cc * 2.0 * 3.14159265358979323846 / p;
            for (int A = 0; A < nA; A++) {
                int ax, ay, az;
                cart_comp(la, A, ax, ay, az);
                for (int B = 0; B < nB; B++) {
                    int bx, by, bz;
                    cart_comp(lb, B, bx, by, bz);
                    const double* ex = Ex + ax * si;
                    const double* ey = Ey + ay * si;
                    const double* ez = Ez + az * si;
                    // overlap factors s(j) = E[i, j, 0] and the second derivative on the ket, one direction each
                    const double sx = ex[bx * sj], sy = ey[by * sj], sz = ez[bz * sj];
                    auto d2 = [&](const double* e, int j) {
                        double v = -2.0 * b * (2 * j + 1) * e[j * sj] + 4.0 * b * b * e[(j + 2) * sj];
                        if (j >= 2) v += (double)(j * (j - 1)) * e[(j - 2) * sj];
                        return v;
                    };
                    const double dx = d2(ex, bx), dy = d2(ey, by), dz = d2(ez, bz);
                    accS[A * nB + B] = fma(cc * rt3, sx * sy * sz, accS[A * nB + B]);
                    accT[A * nB + B] = fma(-0.5 * cc * rt3, dx * sy * sz + sx * dy * sz + sx * sy * dz, accT[A * nB + B]);
                    double s = 0.0;
                    for (int t = 0; t <= ax + bx; t++) {
                        double su = 0.0;
                        for (int u = 0; u <= ay + by; u++) {
                            double sv = 0.0;
                            for (int v = 0; v <= az + bz; v++) sv = fma(ez[bz * sj + v], Rz[ix(t, u, v)], sv);
                            su = fma(ey[by * sj + u], sv, su);
                        }
                        s = fma(ex[bx * sj + t], su, s);
                    }
                    accV[A * nB + B] = fma(vpref, s, accV[A * nB + B]);
                }
            }
        }
    const int fa0 = g.prim.fn_off[sa], fb0 = g.prim.fn_off[sb];
    for (int A = 0; A < nA; A++) {
        int ax, ay, az;
        cart_comp(la, A, ax, ay, az);
        const double fA = comp_factor(la, ax, ay, az);
        for (int B = 0; B < nB; B++) {
            int bx, by, bz;
            cart_comp(lb, B, bx, by, bz);
            const double f = fA * comp_factor(lb, bx, by, bz);
            const int64_t i1 = (fa0 + A) + (int64_t)g.N * (fb0 + B), i2 = (fb0 + B) + (int64_t)g.N * (fa0 + A);
            g.S[i1] = g.S[i2] = f * accS[A * nB + B];
            g.T[i1] = g.T[i2] = f * accT[A * nB + B];
            g.V[i1] = g.V[i2] = f * accV[A * nB + B];
        }
    }
}

}  // namespace jcdf

extern "C" int32_t jcdf_compute_one_electron(jcdf_handle* h, int32_t natoms, const double* charges, const double* centers,
                                             double* S_out, double* T_out, double* V_out, double* E_nuc_out) {
    JC_REQUIRE(h && charges && centers && natoms >= 1, "jcdf_compute_one_electron: bad argument");
    IntegralsState* st = h->ints;
    JC_REQUIRE(st && st->have[0], "jcdf_compute_one_electron: call jcdf_set_basis(which = 0) first");
    Shard& s0 = h->shards[0];
    const int64_t N = h->lay.N, nn = N * N;
    JC_CUDA(cudaSetDevice(s0.dev));
    double *dS = nullptr, *dZ = nullptr, *dC = nullptr;
    JC_CUDA(cudaMalloc(reinterpret_cast<void**>(&dS), sizeof(double) * 3 * nn));
    JC_CUDA(cudaMalloc(reinterpret_cast<void**>(&dZ), sizeof(double) * natoms));
    JC_CUDA(cudaMalloc(reinterpret_cast<void**>(&dC), sizeof(double) * 3 * natoms));
    JC_CUDA(cudaMemcpyAsync(dZ, charges, sizeof(double) * natoms, cudaMemcpyHostToDevice, s0.stream));
    JC_CUDA(cudaMemcpyAsync(dC, centers, sizeof(double) * 3 * natoms, cudaMemcpyHostToDevice, s0.stream));
    OneElectronArgs a;
    a.prim = st->devs[0].basis[0];
    a.boys_table = st->devs[0].boys_table;
    a.charges = dZ; a.nuclei = dC; a.natoms = natoms;
    a.S = dS; a.T = dS + nn; a.V = dS + 2 * nn;
    a.N = (int)N;
    const int ns = a.prim.nshell;
    const int64_t npairs = (int64_t)ns * (ns + 1) / 2;
    int lmax = 0;
    for (int l : st->host[0].l) lmax = std::max(lmax, l);
    const unsigned grid = (unsigned)((npairs + INT_THREADS - 1) / INT_THREADS);
    if (2 * lmax <= 4) k_one_electron<4><<<grid, INT_THREADS, 0, s0.stream>>>(a);
    else k_one_electron<6><<<grid, INT_THREADS, 0, s0.stream>>>(a);
    JC_CUDA(cudaGetLastError());
    if (S_out) JC_CUDA(cudaMemcpyAsync(S_out, a.S, sizeof(double) * nn, cudaMemcpyDeviceToHost, s0.stream));
    if (T_out) JC_CUDA(cudaMemcpyAsync(T_out, a.T, sizeof(double) * nn, cudaMemcpyDeviceToHost, s0.stream));
    if (V_out) JC_CUDA(cudaMemcpyAsync(V_out, a.V, sizeof(double) * nn, cudaMemcpyDeviceToHost, s0.stream));
    JC_CUDA(cudaStreamSynchronize(s0.stream));
    cudaFree(dS); cudaFree(dZ); cudaFree(dC);
    if (E_nuc_out) {  // nuclear repulsion, pairs in a fixed order
        double e = 0.0;
        for (int i = 0; i < natoms; i++)
            for (int j = 0; j < i; j++) {
                const double dx = centers[3 * i] - centers[3 * j], dy = centers[3 * i + 1] - centers[3 * j + 1],
                             dz = centers[3 * i + 2] - centers[3 * j + 2];
                e += charges[i] * charges[j] / std::sqrt(dx * dx + dy * dy + dz * dz);
            }
        *E_nuc_out = e;
    }
    return JCDF_OK;
}
