"""One-process-per-GPU plumbing (torchrun / MPI ranks): which auxiliary rows a rank owns and how the NCCL unique id
reaches every rank.  torch.distributed is used for the rendezvous only — the per-build reduction is the ncclAllReduce
inside jcdf_fock_build (replacing MPI.Allreduce!, DensityFitting.jl:68-71).  Pure host logic: testable with gloo."""
from __future__ import annotations

from .df import balanced_function_shards, static_load_rank_indicies, unique_id


def rank_shards(Q: int, world: int, rank: int, ndev_local: int = 1, aux_shell_nbas=None) -> list[range]:
    """Auxiliary-function ranges of this rank's devices.  With aux_shell_nbas: the reference rule (contiguous auxiliary
    shells, floor(nsh / pieces) each, remainder to the last piece — DynamicLoad.jl:160-203, per device
    GPUDF.jl:1026-1056); without: an even split by function count."""
    total = world * ndev_local
    if aux_shell_nbas is not None:
        return [static_load_rank_indicies(rank * ndev_local + d, total, aux_shell_nbas)[1] for d in range(ndev_local)]
    allr = balanced_function_shards(Q, total)
    return allr[rank * ndev_local:(rank + 1) * ndev_local]


def exchange_unique_id(dist, rank: int, device=None, make_id=unique_id) -> bytes:
    """Rank 0 creates the 128-byte NCCL id (jcdf_comm_unique_id) and broadcasts it over the already-initialised
    torch.distributed group (any backend); every rank returns the same bytes for jcdf_comm_init_rank."""
    import torch

    buf = torch.zeros(128, dtype=torch.uint8, device=device)
    if rank == 0:
        raw = make_id()
        if len(raw) != 128:
            raise ValueError("NCCL unique id must be 128 bytes")
        buf = torch.tensor(list(raw), dtype=torch.uint8, device=device)
    dist.broadcast(buf, 0)
    return bytes(buf.cpu().tolist())
