"""Cartesian Gaussian integrals for the oracle (TEST INFRASTRUCTURE — never shipped, never on the product path).

The reference gets S, T, V, (P|Q) and (P|mn) from libint2 (libint_jll 2.7.0+0, Manifest.toml:756-760 — an un-vendored
binary that is absent from /root/reference) through the CxxWrap shim deps/src/jeri-oei.hpp / jeri-df-tei.hpp:51-95.
This module restates the published McMurchie-Davidson scheme (Hermite expansion coefficients E_t^{ij}, Hermite Coulomb
integrals R_tuv, Boys function) with the reference's conventions, so that oracle/scf_replay.py can regenerate the
reference's logged water cc-pVDZ / cc-pVDZ-RIFIT run (water_ccpvdz_out.log) and pin oracle/df_oracle.py against it:

  * Cartesian (6d/10f) functions in libint/CCA order xx,xy,xz,yy,yz,zz ... (pure=false, jeri-core.hpp:41-62);
  * libint's Shell::renorm(): each contraction normalised to a unit-norm axis-aligned x^l component;
  * JuliaChem's axial normalisation on top (Globals.jl:20-44 tables, applied at EnergyHelpers.jl:320-411 and in the
    DF integral drivers), which makes EVERY Cartesian component unit-normalised:
    factor = sqrt((2l-1)!! / ((2a-1)!!(2b-1)!!(2c-1)!!));
  * 3-centre blocks indexed [aux fn][mu][nu] (ThreeCenterIntegrals.jl:169-183); the 2-centre metric is symmetric
    (the reference fills the lower triangle only, TwoCenterIntegrals.jl:39-41).
"""
from __future__ import annotations

from dataclasses import dataclass
from math import pi, sqrt

import numpy as np
from scipy import special


def cart_components(l: int):
    """libint / CCA standard order: (l,0,0), (l-1,1,0), (l-1,0,1), ... (0,0,l)."""
    return [(i, j, l - i - j) for i in range(l, -1, -1) for j in range(l - i, -1, -1)]


def dfact(n: int) -> float:
    """(n)!! with (-1)!! = 1."""
    r = 1.0
    while n > 1:
        r *= n
        n -= 2
    return r


@dataclass
class Shell:
    l: int
    exps: np.ndarray      # primitive exponents
    coefs: np.ndarray     # contraction coefficients as printed by the reference (raw BSE values)
    center: np.ndarray    # bohr
    atom: int = 0

    @property
    def nbas(self) -> int:
        return (self.l + 1) * (self.l + 2) // 2

    def normalized_coefs(self) -> np.ndarray:
        """libint2 Shell::renorm(): primitive norm of the axis-aligned component, then unit-normalise the contraction."""
        l, a = self.l, np.asarray(self.exps, dtype=np.float64)
        c = np.asarray(self.coefs, dtype=np.float64) * np.sqrt(
            2.0 ** l * (2.0 * a) ** (l + 1.5) / (pi ** 1.5 * dfact(2 * l - 1)))
        g = a[:, None] + a[None, :]
        norm = (dfact(2 * l - 1) * pi ** 1.5 * c[:, None] * c[None, :] / (2.0 ** l * g ** (l + 1.5))).sum()
        return c / sqrt(norm)

    def component_factors(self) -> np.ndarray:
        """JuliaChem's axial normalisation (Globals.jl:20-44): every Cartesian component unit-normalised."""
        l = self.l
        return np.array([sqrt(dfact(2 * l - 1) / (dfact(2 * i - 1) * dfact(2 * j - 1) * dfact(2 * k - 1)))
                         for (i, j, k) in cart_components(l)])


def shell_offsets(shells):
    pos, n = [], 0
    for s in shells:
        pos.append(n)
        n += s.nbas
    return pos, n


# ----------------------------------------------------------------------------------------------------------------------
# Boys function F_n(x), n = 0..nmax, vectorised over x
# ----------------------------------------------------------------------------------------------------------------------
def boys(nmax: int, x: np.ndarray) -> np.ndarray:
    """Returns F[n, ...].  Top order from the incomplete gamma function (series for tiny x), then the stable downward
    recursion F_{n-1} = (2x F_n + exp(-x)) / (2n - 1)."""
    x = np.asarray(x, dtype=np.float64)
    F = np.empty((nmax + 1,) + x.shape)
    small = x < 1e-6
    xs = np.where(small, 1.0, x)
    top = special.gamma(nmax + 0.5) * special.gammainc(nmax + 0.5, xs) / (2.0 * xs ** (nmax + 0.5))
    series = 1.0 / (2 * nmax + 1) - x / (2 * nmax + 3) + x * x / (2.0 * (2 * nmax + 5))
    F[nmax] = np.where(small, series, top)
    ex = np.exp(-x)
    for n in range(nmax, 0, -1):
        F[n - 1] = (2.0 * x * F[n] + ex) / (2 * n - 1)
    return F


# ----------------------------------------------------------------------------------------------------------------------
# Hermite expansion coefficients (one Cartesian direction)
# ----------------------------------------------------------------------------------------------------------------------
def hermite_E(la: int, lb: int, a: float, b: float, Ax: float, Bx: float) -> np.ndarray:
    """E[i, j, t] with x_A^i x_B^j exp(-a x_A^2 - b x_B^2) = sum_t E[i,j,t] Lambda_t(x - P_x; p), p = a + b.
    b = 0 (and Bx = Ax) gives the expansion of a single Gaussian."""
    p = a + b
    Px = (a * Ax + b * Bx) / p
    XPA, XPB = Px - Ax, Px - Bx
    E = np.zeros((la + 1, lb + 1, la + lb + 2))
    E[0, 0, 0] = np.exp(-(a * b / p) * (Ax - Bx) ** 2)
    inv2p = 0.5 / p
    for i in range(la):
        for t in range(i + 2):
            E[i + 1, 0, t] = (inv2p * E[i, 0, t - 1] if t > 0 else 0.0) + XPA * E[i, 0, t] + (t + 1) * E[i, 0, t + 1]
    for i in range(la + 1):
        for j in range(lb):
            for t in range(i + j + 2):
                E[i, j + 1, t] = (inv2p * E[i, j, t - 1] if t > 0 else 0.0) + XPB * E[i, j, t] + (t + 1) * E[i, j, t + 1]
    return E[:, :, :la + lb + 1]


def hermite_R(L: int, p: np.ndarray, RPC: np.ndarray) -> np.ndarray:
    """Hermite Coulomb integrals R[..., t, u, v] (order n = 0), t+u+v <= L, vectorised over the leading axes of p / RPC
    (RPC[..., 3]).  R^n_000 = (-2p)^n F_n(p R^2); R^n_{t+1,u,v} = t R^{n+1}_{t-1,u,v} + X R^{n+1}_{t,u,v}; etc."""
    p = np.asarray(p, dtype=np.float64)
    X, Y, Z = RPC[..., 0], RPC[..., 1], RPC[..., 2]
    F = boys(L, p * (X * X + Y * Y + Z * Z))
    shape = p.shape
    R = np.zeros((L + 1,) + shape + (L + 1, L + 1, L + 1))  # R[n, ..., t, u, v]
    for n in range(L + 1):
        R[n, ..., 0, 0, 0] = (-2.0 * p) ** n * F[n]
    for n in range(L - 1, -1, -1):
        m = L - n  # total order reachable at this n
        for t in range(m + 1):
            for u in range(m + 1 - t):
                for v in range(m + 1 - t - u):
                    if t + u + v == 0:
                        continue
                    if t > 0:
                        val = X * R[n + 1, ..., t - 1, u, v]
                        if t > 1:
                            val = val + (t - 1) * R[n + 1, ..., t - 2, u, v]
                    elif u > 0:
                        val = Y * R[n + 1, ..., t, u - 1, v]
                        if u > 1:
                            val = val + (u - 1) * R[n + 1, ..., t, u - 2, v]
                    else:
                        val = Z * R[n + 1, ..., t, u, v - 1]
                        if v > 1:
                            val = val + (v - 1) * R[n + 1, ..., t, u, v - 2]
                    R[n, ..., t, u, v] = val
    return R[0]


# ----------------------------------------------------------------------------------------------------------------------
# primitive-pair Hermite expansion of a product of two contracted Cartesian shells
# ----------------------------------------------------------------------------------------------------------------------
def _pair_primitives(sa: Shell, sb: Shell, extra_b: int = 0):
    """Yields (ca*cb, a, b, p, P, Ex, Ey, Ez) for every primitive pair; E arrays go up to lb + extra_b in j."""
    ca, cb = sa.normalized_coefs(), sb.normalized_coefs()
    for ia, a in enumerate(sa.exps):
        for ib, b in enumerate(sb.exps):
            p = a + b
            P = (a * sa.center + b * sb.center) / p
            E = [hermite_E(sa.l, sb.l + extra_b, a, b, sa.center[d], sb.center[d]) for d in range(3)]
            yield ca[ia] * cb[ib], a, b, p, P, E[0], E[1], E[2]


def _pair_tensor(sa: Shell, sb: Shell, Ex, Ey, Ez) -> np.ndarray:
    """Eab[A, B, t, u, v] = Ex[ax,bx,t] Ey[ay,by,u] Ez[az,bz,v] times the component normalisation factors."""
    ca, cb = cart_components(sa.l), cart_components(sb.l)
    fa, fb = sa.component_factors(), sb.component_factors()
    T = sa.l + sb.l + 1
    out = np.zeros((len(ca), len(cb), T, T, T))
    for A, (ax, ay, az) in enumerate(ca):
        for B, (bx, by, bz) in enumerate(cb):
            out[A, B] = (fa[A] * fb[B]) * np.einsum("t,u,v->tuv", Ex[ax, bx, :T], Ey[ay, by, :T], Ez[az, bz, :T])
    return out


# ----------------------------------------------------------------------------------------------------------------------
# one-electron integrals
# ----------------------------------------------------------------------------------------------------------------------
def overlap_kinetic(shells):
    """S and T (N x N)."""
    pos, N = shell_offsets(shells)
    S, T = np.zeros((N, N)), np.zeros((N, N))
    for i, sa in enumerate(shells):
        for j, sb in enumerate(shells[:i + 1]):
            ca, cb = cart_components(sa.l), cart_components(sb.l)
            fa, fb = sa.component_factors(), sb.component_factors()
            s_blk, t_blk = np.zeros((len(ca), len(cb))), np.zeros((len(ca), len(cb)))
            for cc, a, b, p, P, Ex, Ey, Ez in _pair_primitives(sa, sb, extra_b=2):
                rt = sqrt(pi / p)
                E = (Ex, Ey, Ez)

                def s1(d, ia, jb):
                    return E[d][ia, jb, 0] * rt if jb >= 0 else 0.0

                def d1(d, ia, jb):  # second derivative on the ket, one direction
                    return jb * (jb - 1) * s1(d, ia, jb - 2) - 2.0 * b * (2 * jb + 1) * s1(d, ia, jb) + 4.0 * b * b * s1(d, ia, jb + 2)

                for A, la_ in enumerate(ca):
                    for B, lb_ in enumerate(cb):
                        sx, sy, sz = (s1(d, la_[d], lb_[d]) for d in range(3))
                        dx, dy, dz = (d1(d, la_[d], lb_[d]) for d in range(3))
                        f = cc * fa[A] * fb[B]
                        s_blk[A, B] += f * sx * sy * sz
                        t_blk[A, B] += -0.5 * f * (dx * sy * sz + sx * dy * sz + sx * sy * dz)
            S[pos[i]:pos[i] + sa.nbas, pos[j]:pos[j] + sb.nbas] = s_blk
            S[pos[j]:pos[j] + sb.nbas, pos[i]:pos[i] + sa.nbas] = s_blk.T
            T[pos[i]:pos[i] + sa.nbas, pos[j]:pos[j] + sb.nbas] = t_blk
            T[pos[j]:pos[j] + sb.nbas, pos[i]:pos[i] + sa.nbas] = t_blk.T
    return S, T


def nuclear_attraction(shells, charges, centers):
    """V (N x N) = - sum_C Z_C (mu| 1/|r - C| |nu)."""
    pos, N = shell_offsets(shells)
    V = np.zeros((N, N))
    Z = np.asarray(charges, dtype=np.float64)
    Cn = np.asarray(centers, dtype=np.float64)
    for i, sa in enumerate(shells):
        for j, sb in enumerate(shells[:i + 1]):
            L = sa.l + sb.l
            blk = np.zeros((sa.nbas, sb.nbas))
            for cc, a, b, p, P, Ex, Ey, Ez in _pair_primitives(sa, sb):
                Eab = _pair_tensor(sa, sb, Ex, Ey, Ez)
                R = hermite_R(L, np.full(len(Z), p), P[None, :] - Cn)  # (ncenters, L+1, L+1, L+1)
                Rz = np.einsum("c,ctuv->tuv", Z, R)
                blk += (-cc * 2.0 * pi / p) * np.einsum("ABtuv,tuv->AB", Eab, Rz)
            V[pos[i]:pos[i] + sa.nbas, pos[j]:pos[j] + sb.nbas] = blk
            V[pos[j]:pos[j] + sb.nbas, pos[i]:pos[i] + sa.nbas] = blk.T
    return V


def nuclear_repulsion(charges, centers) -> float:
    Z, C = np.asarray(charges, dtype=np.float64), np.asarray(centers, dtype=np.float64)
    e = 0.0
    for i in range(len(Z)):
        for j in range(i):
            e += Z[i] * Z[j] / np.linalg.norm(C[i] - C[j])
    return float(e)


# ----------------------------------------------------------------------------------------------------------------------
# two- and three-centre Coulomb integrals over the auxiliary basis
# ----------------------------------------------------------------------------------------------------------------------
class _AuxGroup:
    """All primitives of the auxiliary shells of one angular momentum, flattened, with their single-Gaussian Hermite
    expansions Ec[k, C, tau, nu, phi] (sign (-1)^(tau+nu+phi), contraction coefficient and component factor folded in)."""

    def __init__(self, l: int, shells, shell_ids):
        self.l = l
        self.shell_ids = shell_ids
        comps = cart_components(l)
        exps, cents, owner, Ec = [], [], [], []
        for si in shell_ids:
            s = shells[si]
            cn, fc = s.normalized_coefs(), s.component_factors()
            for ip, c in enumerate(s.exps):
                E1 = hermite_E(l, 0, c, 0.0, 0.0, 0.0)[:, 0, :]  # [i, tau]; centre-independent
                e = np.zeros((len(comps), l + 1, l + 1, l + 1))
                for C, (cx, cy, cz) in enumerate(comps):
                    e[C] = cn[ip] * fc[C] * np.einsum("t,u,v->tuv", E1[cx], E1[cy], E1[cz])
                sign = (-1.0) ** np.add.outer(np.add.outer(np.arange(l + 1), np.arange(l + 1)), np.arange(l + 1))
                Ec.append(e * sign[None])
                exps.append(c)
                cents.append(s.center)
                owner.append(si)
        self.exps = np.array(exps)
        self.centers = np.array(cents)
        self.owner = np.array(owner)
        self.Ec = np.array(Ec)
        self.Ec_nosign = self.Ec * np.where(self.Ec != 0, 1.0, 1.0)  # same array; sign removed where needed below
        self.sign = sign

    def contract_R(self, R: np.ndarray, T: int) -> np.ndarray:
        """G[k, C, t, u, v] = sum_{tau nu phi} Ec[k, C, tau, nu, phi] R[k, t+tau, u+nu, v+phi]  for t,u,v < T."""
        l = self.l
        G = np.zeros((R.shape[0], self.Ec.shape[1], T, T, T))
        for ta in range(l + 1):
            for nu in range(l + 1 - ta):
                for ph in range(l + 1 - ta - nu):
                    G += self.Ec[:, :, ta, nu, ph][:, :, None, None, None] * R[:, None, ta:ta + T, nu:nu + T, ph:ph + T]
        return G


def _aux_groups(aux_shells):
    by_l = {}
    for i, s in enumerate(aux_shells):
        by_l.setdefault(s.l, []).append(i)
    return [_AuxGroup(l, aux_shells, ids) for l, ids in sorted(by_l.items())]


def _three_center_row(args):
    """All blocks (P | shell i, shell j <= i) of one primary shell i: {j: {aux shell: [C, A, B]}}."""
    i, aux_shells, shells = args
    groups = _aux_groups(aux_shells)
    sa = shells[i]
    row = {}
    for j, sb in enumerate(shells[:i + 1]):
        Lab = sa.l + sb.l
        T = Lab + 1
        blocks = {g.l: np.zeros((len(aux_shells), (g.l + 1) * (g.l + 2) // 2, sa.nbas, sb.nbas)) for g in groups}
        for cc, a, b, p, P, Ex, Ey, Ez in _pair_primitives(sa, sb):
            Eab = cc * _pair_tensor(sa, sb, Ex, Ey, Ez)
            for g in groups:
                q = g.exps
                alpha = p * q / (p + q)
                R = hermite_R(Lab + g.l, alpha, P[None, :] - g.centers)
                # pad so that slices [tau:tau+T] exist for every tau <= l
                G = g.contract_R(R, T)
                pref = 2.0 * pi ** 2.5 / (p * q * np.sqrt(p + q))
                val = np.einsum("k,kCtuv,ABtuv->kCAB", pref, G, Eab)
                np.add.at(blocks[g.l], g.owner, val)
        row[j] = {si: blocks[g.l][si] for g in groups for si in g.shell_ids}
    return i, row


def _pool_map(fn, jobs, workers: int):
    """Plain loop, or a fork pool over independent jobs (the arithmetic of every job is unchanged)."""
    if workers <= 1 or len(jobs) < 2:
        return [fn(j) for j in jobs]
    import multiprocessing as mp
    with mp.get_context("fork").Pool(min(workers, len(jobs))) as pool:
        return pool.map(fn, jobs, chunksize=1)


def three_center_integrals(aux_shells, shells, workers: int = 1) -> np.ndarray:
    """(P|mu nu), returned as a C-ordered array T[P, mu, nu] (symmetric in mu, nu).  workers > 1: primary shells are
    dealt to a fork pool (same arithmetic per block)."""
    apos, Q = shell_offsets(aux_shells)
    pos, N = shell_offsets(shells)
    out = np.zeros((Q, N, N))
    order = sorted(range(len(shells)), key=lambda i: -(i + 1) * shells[i].nbas)  # heaviest rows first
    for i, row in _pool_map(_three_center_row, [(i, aux_shells, shells) for i in order], workers):
        sa = shells[i]
        for j, per_aux in row.items():
            sb = shells[j]
            for si, blk in per_aux.items():  # [C, A, B]
                n = aux_shells[si].nbas
                out[apos[si]:apos[si] + n, pos[i]:pos[i] + sa.nbas, pos[j]:pos[j] + sb.nbas] = blk
                out[apos[si]:apos[si] + n, pos[j]:pos[j] + sb.nbas, pos[i]:pos[i] + sa.nbas] = blk.transpose(0, 2, 1)
    return out


def two_center_integrals(aux_shells) -> np.ndarray:
    """(P|Q), full symmetric Q x Q matrix."""
    apos, Q = shell_offsets(aux_shells)
    out = np.zeros((Q, Q))
    groups = _aux_groups(aux_shells)
    for i, sa in enumerate(aux_shells):
        ca_n, fa = sa.normalized_coefs(), sa.component_factors()
        comps = cart_components(sa.l)
        T = sa.l + 1
        for ip, a in enumerate(sa.exps):
            E1 = hermite_E(sa.l, 0, a, 0.0, 0.0, 0.0)[:, 0, :]
            Ea = np.zeros((len(comps), 1, T, T, T))
            for A, (ax, ay, az) in enumerate(comps):
                Ea[A, 0] = ca_n[ip] * fa[A] * np.einsum("t,u,v->tuv", E1[ax], E1[ay], E1[az])
            for g in groups:
                q = g.exps
                alpha = a * q / (a + q)
                R = hermite_R(sa.l + g.l, alpha, sa.center[None, :] - g.centers)
                G = g.contract_R(R, T)
                pref = 2.0 * pi ** 2.5 / (a * q * np.sqrt(a + q))
                val = np.einsum("k,kCtuv,ABtuv->kCA", pref, G, Ea)  # B has length 1
                for k, si in enumerate(g.owner):
                    n = aux_shells[si].nbas
                    out[apos[i]:apos[i] + sa.nbas, apos[si]:apos[si] + n] += val[k].T
    return out


def schwarz_block(sa: Shell, sb: Shell) -> np.ndarray:
    """Full (ab|ab) block of one shell pair, shape (nA, nB, nA, nB) = (mu, nu | mu', nu'), with the axial normalisation
    on all four indices — what compute_eri_block(ish, jsh, ish, jsh) + axial_normalization_factor hand to
    schwarz_screen_itegrals_df (SchwarzScreening.jl:41-44).  flatten_schwarz_block() gives the reference's 1-D order."""
    Lab = sa.l + sb.l
    T = Lab + 1
    prims = [(cc, p, P, _pair_tensor(sa, sb, Ex, Ey, Ez)) for cc, a, b, p, P, Ex, Ey, Ez in _pair_primitives(sa, sb)]
    sign = (-1.0) ** np.add.outer(np.add.outer(np.arange(T), np.arange(T)), np.arange(T))
    out = np.zeros((sa.nbas, sb.nbas, sa.nbas, sb.nbas))
    for cc1, p1, P1, E1 in prims:
        for cc2, p2, P2, E2 in prims:
            alpha = p1 * p2 / (p1 + p2)
            R = hermite_R(2 * Lab, np.array(alpha), (P1 - P2))
            E2s = E2 * sign[None, None]
            G = np.zeros((sa.nbas, sb.nbas, T, T, T))
            for ta in range(T):
                for nu in range(T - ta):
                    for ph in range(T - ta - nu):
                        G += E2s[:, :, ta, nu, ph][:, :, None, None, None] * R[None, None, ta:ta + T, nu:nu + T, ph:ph + T]
            pref = cc1 * cc2 * 2.0 * pi ** 2.5 / (p1 * p2 * sqrt(p1 + p2))
            out += pref * np.einsum("ABtuv,CDtuv->ABCD", E1, G)
    return out


def flatten_schwarz_block(blk: np.ndarray) -> np.ndarray:
    """Index [nu + n_nu*mu + n_nu*n_mu*nu' + n_nu*n_mu*n_nu*mu'] (SchwarzScreening.jl:62)."""
    return np.ascontiguousarray(blk.transpose(2, 3, 0, 1)).ravel()


def _schwarz_row(args):
    i, shells = args
    return i, [flatten_schwarz_block(schwarz_block(shells[i], shells[j])) for j in range(i + 1)]


def schwarz_blocks(shells, workers: int = 1) -> dict:
    """{(ish, jsh): flattened (ab|ab) block} for every shell pair ish >= jsh (input of df_oracle.schwarz_function_mask)."""
    order = sorted(range(len(shells)), key=lambda i: -(i + 1) * shells[i].nbas ** 2)
    out = {}
    for i, row in _pool_map(_schwarz_row, [(i, shells) for i in order], workers):
        for j, blk in enumerate(row):
            out[(i, j)] = blk
    return out
