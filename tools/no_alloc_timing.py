#!/usr/bin/env python
"""Device time of jcdf_fock_build_no_alloc next to the handle-based jcdf_fock_build_device on the same small dense problem
(CUDA events on the launching stream, inputs resident).  Prints one JSON line per shape."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import juliachem_jl_b200 as jc  # noqa: E402
from helpers import random_inputs, reference_layout_rows  # noqa: E402


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=np.float64).ravel(order="F"))).cuda()


def main():
    torch.cuda.set_device(0)
    for N, Q, occ in ((25, 116, 5), (47, 166, 5), (48, 256, 16)):
        B, C, H = random_inputs(N, Q, occ, seed=5)
        dB, dC, dH = dev(B.reshape((Q, N * N), order="F")), dev(C), dev(H)
        dF = torch.empty(N * N, dtype=torch.float64, device="cuda")
        gd = jc.B200Data(N, Q, occ, 1)
        gd.set_mask(jc.ScreeningData(None, N))
        gd.set_B(0, reference_layout_rows(B, range(Q), None))
        dF2 = torch.empty(N * N, dtype=torch.float64, device="cuda")
        st = torch.cuda.ExternalStream(gd.stream(0))
        out = {"N": N, "Q": Q, "occ": occ}
        for name, call, stream in (
                ("no_alloc_ms", lambda: jc.fock_build_no_alloc(N, Q, occ, dB.data_ptr(), Q, dC.data_ptr(), N, dH.data_ptr(), dF.data_ptr()),
                 torch.cuda.default_stream()),
                ("handle_device_ms", lambda: gd.fock_build_device(dC.data_ptr(), N, dH.data_ptr(), dF2.data_ptr()), st)):
            for _ in range(20):
                call()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(200):
                call()
            e1.record(stream)
            torch.cuda.synchronize()
            out[name] = e0.elapsed_time(e1) / 200
        out["max_abs_diff"] = float((dF - dF2).abs().max())
        gd.close()
        print(json.dumps(out))


if __name__ == "__main__":
    main()
