#!/usr/bin/env python
"""Schwarz mask of an S22 complex (6-31G(2df,p), df_sigma = 1e-5) computed by the ORACLE (reference rule,
SchwarzScreening.jl:9-83, tau = sigma^2 / max_P (P|P)) and committed as tests/golden/s22_<n>_mask_sigma1e-5.npy (packed bits), so
bench.py's CPU legs can run without a GPU and the device mask has an independent fixture to equal.
    python tests/golden/make_s22_mask.py [11]"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import df_oracle as orc  # noqa: E402
from oracle import gaussian_integrals as gi  # noqa: E402
from oracle.water_case import shells_from  # noqa: E402
from test_oracle_s22_reference_suite import GOLDEN, oracle_case  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 11
    case = oracle_case(GOLDEN["molecules"][str(n)], GOLDEN["basis"])
    prim = shells_from(case["basis"], case["geometry"])
    aux = shells_from(case["auxiliary_basis"], case["geometry"])
    J = gi.two_center_integrals(aux)
    thr = orc.schwarz_threshold(1e-5, float(np.abs(np.diag(J)).max()))
    pos, N = gi.shell_offsets(prim)
    _, mask = orc.schwarz_function_mask(gi.schwarz_blocks(prim, os.cpu_count() or 1), pos, [s.nbas for s in prim], thr)
    assert np.array_equal(mask, mask.T)
    path = os.path.join(HERE, "s22_%d_mask_sigma1e-5.npy" % n)
    np.save(path, np.packbits(mask, axis=None))
    print("wrote", path, "N =", N, "pairs kept", int(mask.sum()), "survival %.4f" % mask.mean())


if __name__ == "__main__":
    main()
