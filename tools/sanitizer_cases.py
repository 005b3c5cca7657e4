#!/usr/bin/env python
"""The smallest parity case of every kernel family, for compute-sanitizer (one --tool per gpurun call):
    compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitizer_cases.py
Each case builds one Fock matrix through the C-ABI and checks it against the oracle (1e-10), so a sanitizer report can be
read next to a known-correct result.  Families: tile half-transform (MT 2 and 6) with the fused cooperative / the
two-pass Coulomb; per-function half-transform 2 x 4 and 8 x 1 with V from the accumulators, with the density + V pass, and
with the cooperative compact Coulomb; sliced exchange with reduce; graph replay and eager launch."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from helpers import banded_mask, jc, oracle_fock, random_inputs, reference_layout_rows  # noqa: E402

CASES = [
    # name, N, Q, occ, options
    ("tiles_fused_coulomb", 70, 40, 9, {"ht_mode": 1, "coulomb_fused": 1}),
    ("tiles_two_pass_coulomb", 70, 40, 41, {"ht_mode": 1, "coulomb_fused": 0}),
    ("pp_2x4_v_in_ht", 96, 140, 40, {"ht_mode": 2, "coulomb_pp_fused": 0, "coulomb_v_in_ht": 1}),
    ("pp_2x4_v_in_ht_9_warps", 96, 140, 40, {"ht_mode": 2, "coulomb_pp_fused": 0, "coulomb_v_in_ht": 1, "pp_bigreg": 0, "exchange_producer": 0}),
    ("pp_2x4_density_v_pass", 96, 140, 40, {"ht_mode": 2, "coulomb_pp_fused": 0, "coulomb_v_in_ht": 0}),
    ("pp_8x1_v_in_ht", 96, 129, 81, {"ht_mode": 2, "coulomb_pp_fused": 0, "pp_layout": 1}),
    ("pp_coulomb_cooperative", 96, 140, 24, {"ht_mode": 2, "coulomb_pp_fused": 1}),
    ("pp_two_ctas_per_sm_v_in_ht", 96, 140, 40, {"ht_mode": 2, "coulomb_pp_fused": 0, "pp_layout": 2}),
    ("pp_two_ctas_per_sm_v_pass", 96, 129, 81, {"ht_mode": 2, "coulomb_pp_fused": 0, "coulomb_v_in_ht": 0, "pp_layout": 2}),
    # df_exchange_screen: width-14 K blocks, the far corner blocks dropped; F_out is in/out and starts as zeros (iteration 1)
    ("block_screened_exchange", 140, 64, 33, {"ht_mode": 2, "exchange_block_screen": 10}),
    ("sliced_exchange_eager", 140, 64, 33, {"ht_mode": 2, "exchange_slices": 4, "graph": 0}),
    ("dense_no_mask", 64, 32, 8, {}),
]


def main():
    args = sys.argv[1:]
    n_fuzz = int(args[args.index("--fuzz") + 1]) if "--fuzz" in args else 0
    only = [a for i, a in enumerate(args) if a not in ("--fire", "--fuzz") and not (i > 0 and args[i - 1] == "--fuzz")]
    lib = jc.load_library()
    print("library: %s  (bounds-checked build: %s)" % (jc.LIB_PATH, bool(lib.jcdf_checked_selftest(0))), flush=True)
    worst = 0.0
    for name, N, Q, occ, opts in CASES:
        if only and name not in only:
            continue
        mask = None if name == "dense_no_mask" else banded_mask(N, max(8, N // 3), seed=N)
        B, C_occ, H = random_inputs(N, Q, occ, seed=N + occ, mask=mask)
        gd = jc.B200Data(N, Q, occ, 1)
        for k, v in opts.items():
            gd.set_option(k, v)
        gd.set_mask(jc.ScreeningData(mask, N))
        gd.set_B(0, reference_layout_rows(B, range(0, Q), mask))
        blocked = "exchange_block_screen" in opts
        zeros = (lambda: np.zeros((N, N), order="F")) if blocked else (lambda: None)
        F = gd.fock_build(C_occ, H, zeros())[0].copy()
        F2 = gd.fock_build(C_occ, None, zeros())[0]
        if blocked:
            from helpers import orc
            sd = orc.ScreeningData(mask)
            Cfull = np.zeros((N, N), order="F")
            Cfull[:, :occ] = C_occ
            F_ref = orc.df_rhf_fock_build([orc.dense_to_screened(B, sd)], Cfull, occ, H, sd, "screened",
                                          df_exchange_n_blocks=opts["exchange_block_screen"], df_screen_exchange=True)
            assert gd.layout().exchange_blocks_kept < gd.layout().exchange_blocks_total
        else:
            F_ref = oracle_fock(B, C_occ, H, mask)
        err = float(np.abs(F - F_ref).max())
        li = gd.layout()
        fuzz_note = ""
        if n_fuzz:  # schedule fuzzing (checked build): random sleeps at every barrier / copy / flag; results must not move
            worst_f, changed = 0.0, 0
            for seed in range(1, n_fuzz + 1):
                gd.checked_set_fuzz(seed)
                Ff = gd.fock_build(C_occ, None, zeros())[0]
                d = float(np.abs(Ff - F).max())
                worst_f = max(worst_f, d)
                changed += int(not np.array_equal(Ff, F))
            gd.checked_set_fuzz(0)
            fuzz_note = "  fuzz seeds=%d changed=%d max|dF|=%.1e" % (n_fuzz, changed, worst_f)
            # (the tile family's two-pass J sweep adds row slabs with atomics when rows are short: equal to rounding only)
            if worst_f > (1e-12 if name == "tiles_two_pass_coulomb" else 0.0):
                raise SystemExit("%s: the result depends on the schedule (max|dF| = %.3e)%s" % (name, worst_f, fuzz_note))
        gd.close()
        worst = max(worst, err)
        print("%-26s N=%d Q=%d occ=%d launches/build=%d graph_replays=%d  max|F-F_oracle|=%.2e  repeat identical=%s" % (
            name, N, Q, occ, li.kernel_launches_per_build, li.graph_replays, err, bool(np.array_equal(F, F2))) + fuzz_note, flush=True)
    if not worst < 1e-10:
        raise SystemExit("parity failed: %.3e" % worst)
    if "--fire" in sys.argv:  # last: a check that MUST fail (kills the context)
        rc = lib.jcdf_checked_selftest(1)
        print("self-test of the checks: CUDA error %d after a deliberate violation (0 = not a checked build, < 0 = no trap)" % rc, flush=True)
        if rc <= 0:
            raise SystemExit("the bounds checks did not fire")


if __name__ == "__main__":
    main()
