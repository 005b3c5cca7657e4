"""The seam with several PROCESSES (one per GPU), as the reference runs under MPI: every rank calls df_rhf_fock_build_B200 with
its rank, the rank count and the NCCL id (julia/B200DF.jl does the same with MPI.Comm_rank / MPI.Bcast!), shards by the
reference's auxiliary-shell rule, computes its own rows of B on its device (jcdf_compute_B) and receives the full Fock matrix
from the NCCL all-reduce inside the library.  Checked on every rank against the oracle's screened build over the same number
of emulated ranks — on the B rows read back from the devices (1e-10) and on the oracle's own integrals (5e-9: the metric's
conditioning) — and for bit-identity across ranks.  Needs as many GPUs as ranks."""
import json
import os
import subprocess
import sys

import pytest

from helpers import ROOT, jc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [2, 3])
def test_seam_on_several_ranks_matches_the_oracle(world):
    if jc.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + world), os.path.join(ROOT, "tests", "multirank", "seam_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    lines = [json.loads(l) for l in out.stdout.splitlines() if l.startswith("{")]
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert sorted(l["rank"] for l in lines) == list(range(world))
    assert sum(l["q_local"] for l in lines) == 332          # the shards of all ranks cover the auxiliary basis exactly once
    for l in lines:
        assert l["max_abs_err"] < 1e-10 and l["max_abs_err_independent_integrals"] < 5e-9
        assert l["identical_to_rank0"] and l["allreduce_key"]
