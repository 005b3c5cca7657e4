#!/usr/bin/env python
"""Block-screened exchange (option "exchange_block_screen" = the reference's df_exchange_screen, ScreenedDF.jl:385-545) at
the benchmark sizes: for several df_exchange_n_blocks the K blocks the reference would compute, the 128 x 128 exchange tiles
libjcdf leaves out of its work lists, the stage times of a build, and the checks of the result —
  * elements of computed blocks: the same handle's default build (option 0) to 1e-11, and for the real w50 also the committed
    ORACLE fixture of the default build (tests/golden/w50_fock_fixture.json), 1e-10;
  * elements of dropped blocks: exactly F_in + H (what the reference leaves in scf_data.two_electron_fock), bit for bit.
One JSON line per (workload, block count).
    python tools/block_screen_cases.py [workload] [n_blocks ...]     (default: w50 0 10 25 50)"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import juliachem_jl_b200 as jc  # noqa: E402
from juliachem_jl_b200 import rhf, workloads as wl  # noqa: E402


dropped_elements = wl.exchange_dropped_elements


def main():
    args = sys.argv[1:]
    name = args.pop(0) if args and not args[0].lstrip("-").isdigit() else "w50"
    blocks = [int(a) for a in args] or [0, 10, 25, 50]
    if 0 not in blocks:
        blocks = [0] + blocks
    blocks.sort()
    w = wl.WORKLOADS[name]
    N, Q, occ = w.N, w.Q, w.occ
    mol = rhf.run_basis(wl.real_molecule(w), {"basis": w.basis}) if w.real_molecules else None
    C_occ, H = wl.occupied_coefficients(w), wl.core_hamiltonian(w)
    fx_path = os.path.join(ROOT, "tests", "golden", "%s_fock_fixture.json" % w.name)
    fx = json.load(open(fx_path)) if os.path.exists(fx_path) else None
    rng = np.random.default_rng(5)
    F_in = rng.standard_normal((N, N))
    F_in = np.asfortranarray(F_in + F_in.T)
    F_default = None
    for nb in blocks:
        gd = jc.B200Data(N, Q, occ, 1)
        gd.set_option("exchange_block_screen", nb)
        if mol is not None:
            gd.set_basis(0, *mol["primary"].args())
            gd.set_basis(1, *mol["auxiliary"].args())
            gd.compute_two_center(False)
            mask, _ = gd.compute_schwarz_mask(1e-5)
            gd.compute_B()
        else:
            mask = wl.screening_mask(w)
            gd.set_mask(jc.ScreeningData(mask, N))
            gd.fill_synthetic_B(w.seed, w.decay_length)
        li = gd.layout()
        dropped, total, kept = dropped_elements(mask, nb)
        assert (li.exchange_blocks_total, li.exchange_blocks_kept) == (total, kept), (li.exchange_blocks_total, li.exchange_blocks_kept, total, kept)
        ms = []
        F = None
        for _ in range(6):
            F = F_in.copy(order="F")
            gd.fock_build(C_occ, H, F)
            ms.append(gd.last_kernel_ms(0))
        med = {k: float(np.median([m[k] for m in ms[1:]])) for k in ms[0]}
        if nb == 0:
            F_default = F.copy()
        err_default = float(np.abs(F - F_default)[~dropped].max())
        err_fx, n_live = None, None
        if fx is not None:
            idx, val = np.array(fx["index"]), np.array(fx["value"])
            live = ~dropped[idx[:, 0], idx[:, 1]]
            n_live = int(live.sum())
            err_fx = float(np.abs(F[idx[live, 0], idx[live, 1]] - val[live]).max())
            assert err_fx < 1e-10
        exact_dropped = bool(np.array_equal(F[dropped], (F_in + H)[dropped])) if dropped.any() else True
        assert err_default < 1e-11 and exact_dropped and np.array_equal(F, F.T), (err_default, exact_dropped)
        print(json.dumps({"workload": "%s: %s" % (w.name, w.config), "N": N, "Q": Q, "occ": occ, "df_exchange_n_blocks": nb,
                          "K_block_width": int(li.exchange_block_width), "blocks_total": total, "blocks_kept": kept,
                          "elements_dropped_fraction": round(float(dropped.mean()), 4),
                          "exchange_tiles_total": int(li.exchange_tiles_total), "exchange_tiles_skipped": int(li.exchange_tiles_skipped),
                          "exchange_half_tiles_skipped": int(li.exchange_half_tiles_skipped),
                          "stage_ms_median_of_5": {k: round(v, 4) for k, v in med.items()},
                          "max_abs_err_kept_vs_default_build": err_default,
                          "fixture_samples_in_kept_blocks": n_live, "max_abs_err_kept_vs_oracle_fixture": err_fx,
                          "dropped_elements_equal_F_in_plus_H_bitwise": exact_dropped}), flush=True)
        gd.close()


if __name__ == "__main__":
    main()
