"""world_size-2 `gloo` test of the N > 1 host path (CPU): shard ownership, NCCL-id exchange, and the reduction
semantics of the dispatcher (H added by rank 0 only, SUM over ranks == whole; DensityFitting.jl:62-71) with the
oracle standing in for each rank's device work."""
import os
import subprocess
import sys

import numpy as np

from helpers import ROOT

WORKER = r'''
import os, sys
sys.path.insert(0, os.environ["JC_ROOT"]); sys.path.insert(0, os.path.join(os.environ["JC_ROOT"], "tests"))
import numpy as np, torch, torch.distributed as dist
from helpers import random_inputs, oracle_fock, orc
from juliachem_jl_b200 import multi
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
raw = multi.exchange_unique_id(dist, rank, None, make_id=lambda: bytes((7 * i + 3) % 256 for i in range(128)))
assert raw == bytes((7 * i + 3) % 256 for i in range(128))
N, Q, occ = 70, 41, 9
nbas = [1] * 11 + [3] * 6 + [6] * 2
B, C_occ, H = random_inputs(N, Q, occ, seed=5)
mine = multi.rank_shards(Q, world, rank, 1, nbas)[0]
owned = torch.zeros(Q, dtype=torch.int32); owned[mine.start:mine.stop] = 1
dist.all_reduce(owned)
assert bool((owned == 1).all()), "shards must tile [0, Q) exactly once"
Cfull = np.zeros((N, N), order="F"); Cfull[:, :occ] = C_occ
part = orc.df_rhf_fock_build([np.asfortranarray(B[mine.start:mine.stop])], Cfull, occ,
                             H if rank == 0 else np.zeros_like(H), None, "dense")
t = torch.from_numpy(np.ascontiguousarray(part))
dist.all_reduce(t, op=dist.ReduceOp.SUM)
whole = oracle_fock(B, C_occ, H, None)
err = float(np.abs(t.numpy() - whole).max())
assert err < 1e-11, err
even = multi.rank_shards(Q, world, rank, 2)
assert len(even) == 2 and even[0].stop == even[1].start
dist.barrier(); dist.destroy_process_group()
print("rank %d ok %.2e" % (rank, err))
'''


def test_two_rank_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, JC_ROOT=ROOT, MASTER_ADDR="127.0.0.1", MASTER_PORT="29533", WORLD_SIZE="2",
               OMP_NUM_THREADS="2")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r), LOCAL_RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, o
        assert "rank %d ok" % r in o


def test_bench_parses_the_gpu_cpu_affinity_table():
    """bench.py binds every rank to the host cores next to its GPU (nvidia-smi topo -m, tab-separated)."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("bench", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    topo = ("\tGPU0\tGPU1\tNIC0\tCPU Affinity\tNUMA Affinity\tGPU NUMA ID\n"
            "GPU0\t X \tNV18\tPIX\t0-3,8-9\t0\t\tN/A\n"
            "GPU1\tNV18\t X \tSYS\t56-57\t1\t\tN/A\n"
            "NIC0\tPIX\tSYS\t X \t\t\t\t\n\nLegend:\n  X    = Self\n")
    assert bench.parse_cpu_affinity(topo, 0) == {0, 1, 2, 3, 8, 9}
    assert bench.parse_cpu_affinity(topo, 1) == {56, 57}
