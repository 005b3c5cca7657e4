"""Builds a whole DF-RHF problem (S, T, V, (P|Q), (P|mn), Schwarz mask, B) for a molecule given as a golden-fixture
`case` dict (geometry in bohr + shells as the reference printed them) with the oracle's own integrals, and drives the
reference's DEFAULT contraction mode over n MPI ranks emulated as shards (TEST INFRASTRUCTURE — never shipped).

Follows the reference's iteration-1 sequence for contraction_mode "default"/"screened" (ScreenedDF.jl:79-131):
  two-centre integrals (TwoCenterIntegrals.jl:7)  ->  Schwarz mask with tau = sigma^2 / max_P (P|P)
  (SchwarzScreening.jl:9-83)  ->  screened three-centre integrals of the rank's auxiliary shells
  (DynamicLoad.jl:160-203)  ->  B rows = Linv[rank rows, :] T (ScreenedDF.jl:92-190)."""
from __future__ import annotations

import numpy as np

from . import df_oracle as orc
from . import gaussian_integrals as gi
from .water_case import shells_from


def build_problem(case: dict, sigma: float = 1e-5, n_ranks: int = 1, workers: int = 1) -> dict:
    prim = shells_from(case["basis"], case["geometry"])
    aux = shells_from(case["auxiliary_basis"], case["geometry"])
    centers = [a["xyz_bohr"] for a in case["geometry"]]
    S, T = gi.overlap_kinetic(prim)
    V = gi.nuclear_attraction(prim, case["charges"], centers)
    J = gi.two_center_integrals(aux)
    T3 = gi.three_center_integrals(aux, prim, workers)
    Q, N = T3.shape[0], T3.shape[1]
    pos, _ = gi.shell_offsets(prim)
    if sigma > 0:
        thr = orc.schwarz_threshold(sigma, float(np.abs(np.diag(J)).max()))
        _, fn_mask = orc.schwarz_function_mask(gi.schwarz_blocks(prim, workers), pos, [s.nbas for s in prim], thr)
    else:
        fn_mask = np.ones((N, N), dtype=bool, order="F")
    sd = orc.ScreeningData(fn_mask)
    # (Q, N, N) symmetric in the last two -> reference pair order (Q, P)
    T2 = orc.dense_to_screened(np.asfortranarray(T3), sd)
    Linv = orc.metric_inverse_cholesky(np.tril(J))
    nbas_aux = [s.nbas for s in aux]
    ranges = [orc.static_load_rank_indicies(r, n_ranks, nbas_aux)[1] for r in range(n_ranks)]
    B_shards = [orc.calculate_B_rows(Linv, T2, r) for r in ranges]
    return {"S": S, "T": T, "V": V, "H": np.asfortranarray(T + V), "E_nuc": gi.nuclear_repulsion(case["charges"], centers),
            "J_PQ": J, "T3": T3, "sd": sd, "fn_mask": fn_mask, "B_shards": B_shards, "ranges": ranges,
            "nocc": case["n_electrons"] // 2, "aux_shell_nbas": nbas_aux, "N": N, "Q": Q,
            "prim": prim, "aux": aux}
