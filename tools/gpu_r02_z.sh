#!/bin/bash
# round 2, call Z: ncu --set full of the per-build kernels of the FINAL tree (the exchange kernel gained the half-tile test)
mkdir -p gpurun_out
timeout 200 ncu --set full --clock-control none --import-source on -k 'regex:^k_exchange(_ws)?$|k_pp_half_transform|k_pp_coulomb_j' -s 9 -c 3 \
    -o gpurun_out/z_prof python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/z_ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/z_ncu_full.log; ls -la gpurun_out/z_prof.ncu-rep
