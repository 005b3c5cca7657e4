#!/usr/bin/env python
"""bench.py — DF-RHF Fock builds/sec of libjcdf on B200 (BASELINE.json's metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload w50] [--impl b200|reference]

One "step" = one Fock build  F = H + 2J[D] - K[D]  over the resident, Q-sharded, tile-packed B tensor
(jcdf_fock_build*: density pack, half-transform W, exchange K, Coulomb V/J, epilogue, NCCL all-reduce).  One-off setup
(mask, layout, B fill) is outside the timed region.

  value     builds/s with C_occ, H and F resident in HBM (jcdf_fock_build_device), CUDA events on the library's stream,
            barrier + synchronize on both sides, max over ranks.
  e2e       the same through the reference-facing C-ABI call jcdf_fock_build with pinned HOST buffers: C_occ H2D and
            F D2H inside the timed region, every step.
  roofline  dominant kernel of the step, algorithmic FLOPs / its CUDA-event duration, against the FP64
            DGEMM peak measured live in this run (cuBLAS through torch.matmul, 8192^3).
  cpu_baseline  oracle port of the reference's threaded screened CPU algorithm on a bounded Q-slab of the same
            workload, timed on this box's host cores (rank 0, N=1 only).

N > 1: launched by torchrun, one process per GPU; rank r owns a contiguous Q shard (DynamicLoad.jl:160-203), the
partial Fock matrices are summed by ncclAllReduce inside jcdf_fock_build.  Total work is fixed -> "strong" scaling.
--impl reference: the reference's CPU algorithm (oracle port; Julia/libint are not installable here)
on rank 0's host cores for the same metric and config.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "DF-RHF Fock builds/sec"
UNIT = "builds/s"
# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture (not measured live:
# a number taken under a profiler is never a bench value, and DRAM counters need the profiler)
NCU_TRAFFIC_SOURCE = "profiles/r01_ncu_w50_full_k.txt (ncu --set full --clock-control none, per launch, bytes)"
NCU_TRAFFIC = {("w50", 1, "exchange"): 58.177695e9 + 0.503058e9,
               ("w50", 1, "half_transform"): 12.975133e9 + 12.111652e9}
FP64_MICROBENCH_PEAK = 37.1  # TFLOP/s, DMMA issue peak measured by tools/microbench_fp64 (profiles/r01_microbench_fp64.log)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="w50")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--opt", action="append", default=[], help="library option name=value (jcdf_set_option)")
    ap.add_argument("--cpu-seconds", type=float, default=4.0, help="target seconds per CPU sample build")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md recipe)
# ----------------------------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(prefix="clocks_", suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            sm_sorted = sorted(sm)
            out.update({"sm_mhz": sm_sorted[len(sm) // 2], "sm_max_mhz": max(mx), "power_w_max": max(pw),
                        "reasons": sorted(reasons), "samples": len(sm)})
        return out


# ----------------------------------------------------------------------------------------------------------------------
# CPU arm: oracle port of the reference's threaded screened algorithm on a bounded Q-slab
# ----------------------------------------------------------------------------------------------------------------------
def cpu_reference_builds_per_sec(w, target_seconds: float, reps: int, warmup: int):
    """Returns (builds/s extrapolated to the full workload, dict describing the sample).  The Fock build is a plain sum
    over auxiliary rows (every stage is linear in Q), so a slab of Qs rows costs Qs/Q of a build."""
    import numpy as np
    from threadpoolctl import threadpool_limits

    from juliachem_jl_b200 import workloads as wl
    from oracle import df_oracle as orc  # bench.py's CPU legs are allowed to execute the oracle

    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        pass
    N, Q, occ = w.N, w.Q, w.occ
    mask = wl.screening_mask(w)
    if mask is None:
        mask = np.asfortranarray(np.ones((N, N), dtype=bool))
    sd = orc.ScreeningData(mask)
    P = sd.screened_indices_count
    # calibrate: DGEMM rate of this host
    with threadpool_limits(limits=cores, user_api="blas"):
        a = np.asfortranarray(np.random.default_rng(0).standard_normal((2048, 2048)))
        (a @ a)
        dt = 1e9
        for _ in range(3):
            t0 = time.perf_counter(); (a @ a); dt = min(dt, time.perf_counter() - t0)
    gflops = 2 * 2048 ** 3 / dt / 1e9
    flops_per_row = 2.0 * occ * P + 1.0 * occ * N * (N + 1) + 4.0 * P
    qs = int(target_seconds * gflops * 1e9 * 0.5 / flops_per_row)
    qs = max(16, min(Q, qs, int(6e9 / (8.0 * P))))
    rng = np.random.default_rng(w.seed)
    B = np.empty((qs, P), dtype=np.float64, order="F")
    blk = rng.standard_normal((min(qs, 32), P)) / np.sqrt(Q)
    for r0 in range(0, qs, blk.shape[0]):  # timing does not depend on the values; repeat a 32-row slab
        n = min(blk.shape[0], qs - r0)
        B[r0:r0 + n, :] = blk[:n]
    C_occ = wl.occupied_coefficients(w)
    best, stages = None, {}
    for i in range(warmup + reps):
        tm = {}
        t0 = time.perf_counter()
        orc.fock_build_screened_rank_threaded(B, C_occ, sd, cores, 0, tm)
        dt = time.perf_counter() - t0
        if i >= warmup and (best is None or dt < best):
            best, stages = dt, tm
    value = 1.0 / (best * Q / qs)
    info = {"cores": cores, "kind": "port",
            "sample": "Q-slab of %d of %d auxiliary rows of %s (N=%d occ=%d P=%d), reference screened algorithm "
                      "(threads over p / K blocks, OpenBLAS), best of %d after %d warm-up; %.2f s per slab, "
                      "extrapolated linearly in Q" % (qs, Q, w.name, N, occ, P, reps, warmup, best),
            "host_dgemm_gflops": round(gflops, 1),
            "stage_seconds_sample": {k: round(v, 4) for k, v in stages.items()}}
    return value, info


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from juliachem_jl_b200 import workloads as wl
    w = wl.WORKLOADS[args.workload]
    t0 = time.perf_counter()
    value, info = cpu_reference_builds_per_sec(w, args.cpu_seconds, max(1, min(args.steps, 3)), max(1, min(args.warmup, 1)))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 / value, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": data_label(w),
        "config": workload_config(w, args.gpus),
        "cpu_baseline": dict(info, value=value, unit=UNIT),
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_seconds": round(time.perf_counter() - t0, 1),
    }
    print(json.dumps(line))
    return 0


def data_label(w):
    if w.real_molecules:
        return ("real: B = L^-1 (P|mn) from cc-pVDZ / cc-pVDZ-RIFIT integrals of the w50 geometry computed on the device, "
                "real Schwarz mask (df_sigma 1e-5); synthetic orthonormal C_occ and symmetric H (timing is value-independent)")
    return "synthetic"


def workload_config(w, n_gpus):
    return {"workload": "%s: %s" % (w.name, w.config), "N": w.N, "Q": w.Q, "occ": w.occ,
            "sharding": "Q over %d GPU(s), one process per GPU" % n_gpus,
            "l2": "inputs larger than L2 (packed B shard streamed from HBM every build)"}


# ----------------------------------------------------------------------------------------------------------------------
# B200 arm
# ----------------------------------------------------------------------------------------------------------------------
def run_b200(args):
    import numpy as np
    import torch

    import juliachem_jl_b200 as jc
    from juliachem_jl_b200 import workloads as wl

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the B200 arm has no CPU fallback")
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        ndev_local, dev_ids, n_total = 1, [local_rank], world
    else:
        # plain `python bench.py --gpus N`: one process driving N devices through the same library
        ndev_local, dev_ids, n_total = args.gpus, list(range(args.gpus)), args.gpus
        torch.cuda.set_device(0)
    dev0 = torch.device("cuda", dev_ids[0])

    w = wl.WORKLOADS[args.workload]
    N, Q, occ = w.N, w.Q, w.occ
    gd = jc.B200Data(N, Q, occ, ndev_local, dev_ids)
    from juliachem_jl_b200 import multi
    my_shards = multi.rank_shards(Q, world if world > 1 else 1, rank, ndev_local)
    gd.set_shards(my_shards)
    for kv in args.opt:
        k, v = kv.split("=")
        gd.set_option(k, int(v))
    t_fill = time.perf_counter()
    if w.real_molecules:
        # real integrals: basis -> (P|Q) -> Schwarz mask (reference rule, df_sigma 1e-5) -> B = L^-1 (P|mn), all on the
        # device; every rank recomputes what it needs (jcdf_compute_B), so there is no setup exchange between ranks
        from juliachem_jl_b200 import rhf
        mol = rhf.run_basis(wl.real_molecule(w), {"basis": "cc-pVDZ"})
        assert (mol["primary"].n_functions, mol["auxiliary"].n_functions, mol["n_electrons"] // 2) == (N, Q, occ)
        gd.set_basis(0, *mol["primary"].args())
        gd.set_basis(1, *mol["auxiliary"].args())
        gd.compute_two_center(False)
        mask_dev, _ = gd.compute_schwarz_mask(1e-5)
        if not np.array_equal(mask_dev, wl.real_mask(w)):
            raise SystemExit("bench.py: device Schwarz mask differs from the committed fixture")
        if world > 1:
            gd.comm_init_rank(world, rank, multi.exchange_unique_id(dist, rank, dev0))
        gd.compute_B()
    else:
        mask = wl.screening_mask(w)
        gd.set_mask(jc.ScreeningData(mask, N))
        if world > 1:
            gd.comm_init_rank(world, rank, multi.exchange_unique_id(dist, rank, dev0))
        gd.fill_synthetic_B(w.seed, w.decay_length)
    t_fill = time.perf_counter() - t_fill
    li = gd.layout()

    C_host = wl.occupied_coefficients(w)
    H_host = wl.core_hamiltonian(w)
    # device-resident operands (column-major N x occ == row-major (occ, N))
    C_dev = torch.from_numpy(np.ascontiguousarray(C_host.T)).to(dev0)
    H_dev = torch.from_numpy(np.ascontiguousarray(H_host.T)).to(dev0)
    F_dev = torch.empty((N, N), dtype=torch.float64, device=dev0)
    ext = torch.cuda.ExternalStream(gd.stream(0), device=dev0)

    def barrier():
        if dist is not None:
            dist.barrier()
        for d in dev_ids:
            torch.cuda.synchronize(d)

    def maxreduce(x: float) -> float:
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev0)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(step_fn, steps):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record(ext)
        for _ in range(steps):
            step_fn()
        ev1.record(ext)
        barrier()
        return maxreduce(ev0.elapsed_time(ev1))

    stage_acc = {}

    def step_device():
        gd.fock_build_device(C_dev.data_ptr(), N, H_dev.data_ptr(), F_dev.data_ptr())
        for k, v in gd.last_kernel_ms(0).items():
            stage_acc[k] = stage_acc.get(k, 0.0) + v

    for _ in range(max(args.warmup, 3)):
        step_device()
    stage_acc.clear()
    sampler = ClockSampler(dev_ids[0])
    if rank == 0:
        sampler.start()
    ms_total = timed(step_device, args.steps)
    clocks = sampler.stop() if rank == 0 else {}
    ms_per_step = ms_total / args.steps
    value = 1e3 / ms_per_step
    stage_ms = {k: v / args.steps for k, v in stage_acc.items()}
    F_chk = F_dev.cpu().numpy()
    if not (np.isfinite(F_chk).all() and np.array_equal(F_chk, F_chk.T)):
        raise SystemExit("bench.py: Fock matrix is not finite/symmetric")

    # ---- end to end through the host-pointer C-ABI call ----------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        C_pin = torch.from_numpy(np.ascontiguousarray(C_host.T)).pin_memory()
        F_pin = torch.empty((N, N), dtype=torch.float64).pin_memory()
        C_np = C_pin.numpy().T  # column-major N x occ view of the pinned buffer
        F_np = F_pin.numpy().T
        gd.fock_build(C_np, H_host, F_np)  # H uploaded once; stays resident (constant over an SCF)

        def step_host():
            gd.fock_build(C_np, None, F_np)

        for _ in range(2):
            step_host()
        ms_e2e = timed(step_host, args.steps) / args.steps
        if not np.allclose(F_np, F_chk, rtol=0, atol=1e-9):
            raise SystemExit("bench.py: host-path Fock differs from device-path Fock")
        e2e = {"value": 1e3 / ms_e2e, "unit": UNIT, "h2d_bytes_per_step": int(C_pin.numel() * 8) * n_total,
               "d2h_bytes_per_step": int(F_pin.numel() * 8) * (world if world > 1 else 1), "ms_per_step": ms_e2e}

    # ---- roofline of the dominant kernels (rank 0's shard) -----------------------------------------------------------
    q_local0 = len(my_shards[0])
    work = wl.algorithmic_work(N, q_local0, occ, li.pairs_kept, li.pairs_tri)
    peak_tf, peak_src = None, None
    if rank == 0:
        n = 8192
        a = torch.randn((n, n), dtype=torch.float64, device=dev0)
        b = torch.randn((n, n), dtype=torch.float64, device=dev0)
        torch.matmul(a, b)
        best = 1e9
        for _ in range(4):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); torch.matmul(a, b); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        peak_tf = 2.0 * n ** 3 / (best * 1e-3) / 1e12
        peak_src = "cuBLAS DGEMM 8192^3 via torch.matmul, best of 4, measured in this run (MEASURED_PEAKS.json has no FP64 entry)"
        del a, b
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (of fallback)"

    def tf(flops, ms):
        return flops / (ms * 1e-3) / 1e12 if ms > 0 else 0.0

    kern = {
        "half_transform": {"bound": "tensor", "ms": stage_ms.get("W", 0.0), "flops": work["w_flops"],
                           "achieved": tf(work["w_flops"], stage_ms.get("W", 0.0)), "unit": "TFLOP/s"},
        "exchange": {"bound": "tensor", "ms": stage_ms.get("K", 0.0), "flops": work["k_flops"],
                     "achieved": tf(work["k_flops"], stage_ms.get("K", 0.0)), "unit": "TFLOP/s"},
        "coulomb": {"bound": "hbm", "ms": stage_ms.get("V", 0.0) + stage_ms.get("J", 0.0), "bytes": work["j_bytes"],
                    "achieved": work["j_bytes"] / ((stage_ms.get("V", 0.0) + stage_ms.get("J", 0.0)) * 1e-3 + 1e-30) / 1e9,
                    "unit": "GB/s", "peak": hbm_peak, "peak_source": hbm_src},
    }
    kern["coulomb"]["frac"] = kern["coulomb"]["achieved"] / hbm_peak
    dom = max(("half_transform", "exchange", "coulomb"), key=lambda k: kern[k]["ms"])
    roofline = None
    if rank == 0:
        for k in ("half_transform", "exchange"):
            kern[k]["peak"] = peak_tf
            kern[k]["frac"] = kern[k]["achieved"] / peak_tf
        d = kern[dom]
        roofline = {"kernel": dom, "bound": d["bound"], "achieved": d["achieved"], "peak": d["peak"], "unit": d["unit"],
                    "frac": d["frac"], "traffic": NCU_TRAFFIC.get((args.workload, n_total, dom)),
                    "traffic_source": NCU_TRAFFIC_SOURCE if (args.workload, n_total, dom) in NCU_TRAFFIC else None,
                    "peak_source": peak_src if d["bound"] == "tensor" else hbm_src,
                    "fp64_dmma_issue_peak_tflops": FP64_MICROBENCH_PEAK,
                    "jk_tensor_frac": tf(work["w_flops"] + work["k_flops"],
                                         stage_ms.get("W", 0) + stage_ms.get("K", 0)) / peak_tf,
                    "build_frac_of_fp64_peak": tf(work["w_flops"] + work["k_flops"] + work["j_flops"], ms_per_step) / peak_tf,
                    "kernels": kern}

    gd.close()
    del F_dev, C_dev, H_dev
    torch.cuda.empty_cache()

    cpu = None
    if rank == 0 and n_total == 1 and not args.no_cpu_baseline:
        v, info = cpu_reference_builds_per_sec(w, args.cpu_seconds, 2, 1)
        cpu = dict(info, value=v, unit=UNIT)

    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return 0
    cfg = workload_config(w, n_total)
    cfg.update({"pairs_kept": li.pairs_kept, "pairs_tri": li.pairs_tri, "tiles_kept": li.tiles_kept,
                "tiles_total": li.tiles_total, "mask_survival": li.pairs_kept / float(N * N),
                "packed_B_bytes_per_gpu": int(li.bytes_B // max(1, ndev_local)), "fill_seconds": round(t_fill, 2)})
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_total, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": data_label(w), "config": cfg, "e2e": e2e, "options": args.opt,
        "gpu_launches": int(li.kernel_launches_per_build) * args.steps * n_total,
        "gpu_launches_per_build_per_gpu": int(li.kernel_launches_per_build),
        "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
        "stage_ms": {k: round(v, 4) for k, v in stage_ms.items()},
    }
    print(json.dumps(line))
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
