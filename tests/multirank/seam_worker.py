"""One rank of a multi-process run of the seam (launched by tests/test_gpu_multirank_seam.py through torch.distributed.run, one
process per GPU): df_rhf_fock_build_B200 with rank / n_ranks / the NCCL id, device-integral route — what julia/B200DF.jl does
under MPI (DensityFitting.jl:27-29, 68-71).  Every rank builds the Fock matrix of the S22 water dimer's hcore-guess orbitals,
checks it against the oracle (all ranks must hold the full, identical matrix) and prints one JSON line."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist

    import juliachem_jl_b200 as jc
    from juliachem_jl_b200 import multi, rhf
    from oracle import df_oracle as orc
    from oracle import molecule_case as mc
    from oracle import scf_replay as sr
    from test_oracle_s22_reference_suite import GOLDEN, oracle_case

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    m = GOLDEN["molecules"]["2"]
    mol = rhf.run_basis(m, {"basis": GOLDEN["basis"]})

    class Basis(jc.IntegralSource):
        N, Q, n_electrons = mol["primary"].n_functions, mol["auxiliary"].n_functions, mol["n_electrons"]

        def aux_shell_nbas(self):
            return mol["auxiliary"].nbas

        def device_basis(self):
            return mol["primary"].args(), mol["auxiliary"].args()

    # the oracle on this rank's host: same problem, the screened algorithm over `world` shell-sharded ranks
    p = mc.build_problem(oracle_case(m, GOLDEN["basis"]), 1e-5, world, 2)
    H, nocc, sd = p["H"], p["nocc"], p["sd"]
    X = sr.orthogonalizer(p["S"])
    _, _, C, _ = orc.iteration(np.array(H), H, X, nocc)
    F_ref = orc.df_rhf_fock_build(p["B_shards"], C, nocc, H, sd, "screened")

    class OneDevice(jc.B200Data):  # this process drives the GPU torchrun gave it
        def __init__(self, N, Q, occ, ndev, ids=None):
            super().__init__(N, Q, occ, 1, [local])

    import juliachem_jl_b200.df as df
    df.B200Data = OneDevice
    comm_id = multi.exchange_unique_id(dist, rank, torch.device("cuda", local))
    scf_data, timing = jc.SCFData(), jc.JCTiming()
    opts = jc.SCFOptions(num_devices=1, df_use_adaptive=False)
    F = df.df_rhf_fock_build_B200(scf_data, Basis(), np.asfortranarray(C[:, :nocc]), 1, opts, H, timing, rank=rank, n_ranks=world,
                                  comm_id=comm_id)
    # (1) same inputs: every rank's rows of B are read back from its device, gathered, and contracted by the oracle as `world`
    #     emulated ranks — the 1e-10 bar; (2) independent inputs: F_ref came from the ORACLE's own integrals, and the metric of
    #     cc-pVTZ-JKFIT is ill-conditioned enough (1e9 .. 1e10) that integrals equal to 1e-13 relative move F by a few 1e-10
    gd = scf_data.gpu_data
    P = sd.screened_indices_count
    mine = gd.get_B_rows(0, 0, int(gd.layout().q_local), P)
    shards = [None] * world
    dist.all_gather_object(shards, mine)
    F_same = orc.df_rhf_fock_build([np.asfortranarray(b) for b in shards], C, nocc, H, sd, "screened")
    err = float(np.abs(F - F_same).max())
    err_independent = float(np.abs(F - F_ref).max())
    t = torch.from_numpy(np.ascontiguousarray(F)).cuda()
    t0 = t.clone()
    dist.broadcast(t0, 0)
    same = bool(torch.equal(t, t0))
    print(json.dumps({"rank": rank, "world": world, "q_local": int(scf_data.gpu_data.layout().q_local), "max_abs_err": err,
                      "max_abs_err_independent_integrals": err_independent, "identical_to_rank0": same, "allreduce_key": jc.JCTiming.key("fock_MPI_time", 1) in timing.timings}), flush=True)
    scf_data.gpu_data.close()
    dist.barrier()
    dist.destroy_process_group()
    if not (err < 1e-10 and err_independent < 5e-9 and same):
        sys.exit(1)


if __name__ == "__main__":
    main()
