// Device helpers of the Gaussian-integral front-end: McMurchie-Davidson scheme over Cartesian Gaussians
// in the reference's conventions — libint/CCA component order, libint Shell::renorm() contraction normalisation
// (deps/src/jeri-core.hpp:41-62) and JuliaChem's per-component axial factors (src/shared/Globals.jl:20-44, applied at
// EnergyHelpers.jl:320-411), so every Cartesian component is unit-normalised.
//
//   [ab|        Hermite expansion of a primitive product: x_A^i x_B^j e^{-a x_A^2 - b x_B^2} = sum_t E^{ij}_t Lambda_t
//   R_tuv       Hermite Coulomb integrals from the Boys function, built in place (one array of (L+1)(L+2)(L+3)/6)
//   Boys F_n    9-term Taylor interpolation of F_L on a 1/8 grid + stable downward recursion (x < 40), closed form of
//               F_0 + upward recursion beyond
//
// Angular momenta are run-time values (each launch is uniform in (la, lb, lc), so warps do not diverge); the template
// parameters only bound the per-thread scratch arrays.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace jcdf {

constexpr int BOYS_TAYLOR = 9;            // terms k = 0..8
constexpr int BOYS_MAX_ORDER = 16;        // highest F_n ever requested (4 * l_max for (ff|ff))
constexpr int BOYS_ROW = BOYS_MAX_ORDER + BOYS_TAYLOR;  // table rows hold F_0 .. F_24
constexpr double BOYS_STEP = 0.125;
constexpr double BOYS_XMAX = 40.0;
constexpr int BOYS_GRID = 321;            // x = 0, 1/8, ..., 40

// Flat description of a basis set on the device (one for the primary, one for the auxiliary basis).
struct BasisDev {
    int nshell = 0;
    int nfn = 0;
    const int* l = nullptr;         // angular momentum per shell
    const int* nprim = nullptr;
    const int* prim_off = nullptr;  // first primitive of the shell in exps / coefs
    const int* fn_off = nullptr;    // first Cartesian function of the shell
    const double* xyz = nullptr;    // [3 * nshell] centres, bohr
    const double* exps = nullptr;
    const double* coefs = nullptr;  // libint-renormalised contraction coefficients
};

__host__ __device__ __forceinline__ int ncart(int l) { return (l + 1) * (l + 2) / 2; }
__host__ __device__ __forceinline__ int nherm(int L) { return (L + 1) * (L + 2) * (L + 3) / 6; }

// c-th Cartesian component of a shell of angular momentum l in libint/CCA order: (l,0,0), (l-1,1,0), (l-1,0,1), ...
__device__ __forceinline__ void cart_comp(int l, int c, int& i, int& j, int& k) {
    int row = 0;  // row r holds i = l - r and r + 1 components
    while (c >= row + 1) { c -= row + 1; row++; }
    i = l - row;
    j = row - c;
    k = c;
}
__device__ __forceinline__ double dfact_odd(int n) {  // (2n-1)!!
    double r = 1.0;
    for (int m = 2 * n - 1; m > 1; m -= 2) r *= (double)m;
    return r;
}
// JuliaChem's axial normalisation factor of component (i,j,k) of an l shell (Globals.jl:20-44)
__device__ __forceinline__ double comp_factor(int l, int i, int j, int k) {
    return sqrt(dfact_odd(l) / (dfact_odd(i) * dfact_odd(j) * dfact_odd(k)));
}

// Index of (t,u,v), t+u+v <= L, in the packed Hermite array of total order L: off[t] + u*(M+1) - u(u-1)/2 + v, M = L-t
struct HermIndex {
    int L;
    int off[BOYS_MAX_ORDER + 2];
    __device__ __forceinline__ void init(int L_) {
        L = L_;
        int o = 0;
        for (int t = 0; t <= L; t++) {
            off[t] = o;
            const int M = L - t;
            o += (M + 1) * (M + 2) / 2;
        }
    }
    __device__ __forceinline__ int operator()(int t, int u, int v) const {
        const int M = L - t;
        return off[t] + u * (M + 1) - (u * (u - 1)) / 2 + v;
    }
};

// F_n(x), n = 0..L
__device__ __forceinline__ void boys(int L, double x, const double* __restrict__ table, double* __restrict__ F) {
    if (x < BOYS_XMAX) {
        const int g = (int)(x * (1.0 / BOYS_STEP) + 0.5);
        const double d = g * BOYS_STEP - x;  // x0 - x
        const double* row = table + (size_t)g * BOYS_ROW + L;
        // F_L(x) = sum_k F_{L+k}(x0) (x0 - x)^k / k!
        double s = row[BOYS_TAYLOR - 1] * (1.0 / 40320.0);
        s = fma(s, d, row[7] * (1.0 / 5040.0));
        s = fma(s, d, row[6] * (1.0 / 720.0));
        s = fma(s, d, row[5] * (1.0 / 120.0));
        s = fma(s, d, row[4] * (1.0 / 24.0));
        s = fma(s, d, row[3] * (1.0 / 6.0));
        s = fma(s, d, row[2] * 0.5);
        s = fma(s, d, row[1]);
        s = fma(s, d, row[0]);
        F[L] = s;
        const double ex = exp(-x), x2 = 2.0 * x;
        for (int n = L; n > 0; n--) F[n - 1] = fma(x2, F[n], ex) / (double)(2 * n - 1);
    } else {
        const double ex = exp(-x), inv2x = 0.5 / x;
        F[0] = 0.5 * sqrt(3.14159265358979323846 / x) * erf(sqrt(x));
        for (int n = 0; n < L; n++) F[n + 1] = ((2 * n + 1) * F[n] - ex) * inv2x;
    }
}

// Hermite Coulomb integrals R_tuv (order 0) for t+u+v <= L at exponent alpha and separation (X,Y,Z), built in place:
// after stage n the array holds R^n for all orders <= L - n  (R^n_000 = (-2 alpha)^n F_n;
// R^n_{t+1,u,v} = t R^{n+1}_{t-1,u,v} + X R^{n+1}_{t,u,v}, likewise in u and v).
__device__ __forceinline__ void hermite_R(int L, double alpha, double X, double Y, double Z, const double* __restrict__ table,
                                          const HermIndex& ix, double* __restrict__ R) {
    double F[BOYS_MAX_ORDER + 1];
    boys(L, alpha * (X * X + Y * Y + Z * Z), table, F);
    double pw = 1.0;
    for (int n = 0; n <= L; n++) { F[n] *= pw; pw *= -2.0 * alpha; }
    for (int n = L; n >= 0; n--) {
        for (int m = L - n; m >= 1; m--) {
            for (int t = m; t >= 0; t--)
                for (int u = m - t; u >= 0; u--) {
                    const int v = m - t - u;
                    double val;
                    if (t > 0) {
                        val = X * R[ix(t - 1, u, v)];
                        if (t > 1) val = fma((double)(t - 1), R[ix(t - 2, u, v)], val);
                    } else if (u > 0) {
                        val = Y * R[ix(t, u - 1, v)];
                        if (u > 1) val = fma((double)(u - 1), R[ix(t, u - 2, v)], val);
                    } else {
                        val = Z * R[ix(t, u, v - 1)];
                        if (v > 1) val = fma((double)(v - 1), R[ix(t, u, v - 2)], val);
                    }
                    R[ix(t, u, v)] = val;
                }
        }
        R[0] = F[n];
    }
}

// One Cartesian direction of the Hermite expansion: E[(i*(lb+1) + j)*(la+lb+1) + t], 0 <= t <= i + j, including the
// Gaussian product factor exp(-a b / p (Ax - Bx)^2) in E[0,0,0].  b = 0 gives the expansion of a single Gaussian.
__device__ __forceinline__ void hermite_E(int la, int lb, double a, double b, double Ax, double Bx, double* __restrict__ E) {
    const double p = a + b;
    const double Px = (a * Ax + b * Bx) / p;
    const double XPA = Px - Ax, XPB = Px - Bx, inv2p = 0.5 / p;
    const int T = la + lb + 1, sj = T, si = (lb + 1) * T;
    for (int k = 0; k < (la + 1) * si; k++) E[k] = 0.0;
    E[0] = exp(-(a * b / p) * (Ax - Bx) * (Ax - Bx));
    for (int i = 0; i < la; i++)
        for (int t = 0; t <= i + 1; t++) {
            double v = XPA * E[i * si + t];
            if (t > 0) v = fma(inv2p, E[i * si + t - 1], v);
            if (t + 1 <= i) v = fma((double)(t + 1), E[i * si + t + 1], v);
            E[(i + 1) * si + t] = v;
        }
    for (int i = 0; i <= la; i++)
        for (int j = 0; j < lb; j++)
            for (int t = 0; t <= i + j + 1; t++) {
                double v = XPB * E[i * si + j * sj + t];
                if (t > 0) v = fma(inv2p, E[i * si + j * sj + t - 1], v);
                if (t + 1 <= i + j) v = fma((double)(t + 1), E[i * si + j * sj + t + 1], v);
                E[i * si + (j + 1) * sj + t] = v;
            }
}

}  // namespace jcdf
