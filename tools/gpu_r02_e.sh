#!/bin/bash
# round 2, call E: final evidence of the benched kernels — bench line (with CPU leg), reference arm, ncu launch list of the
# same command, ncu --set full of the per-build kernels
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/e_bench.json 2> gpurun_out/e_bench.err; echo "bench rc=$?"; tail -c 400 gpurun_out/e_bench.json
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/e_reference.json 2> gpurun_out/e_reference.err; echo "reference rc=$?"; head -c 300 gpurun_out/e_reference.json
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/e_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/e_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/e_ncu_list.log 2>&1
echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:^k_exchange(_ws)?$|k_pp_half_transform|k_pp_coulomb_j|k_pp_vitem_reduce|k_exchange_reduce|k_fock_epilogue' -s 18 -c 6 \
    -o gpurun_out/e_prof python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/e_ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/e_ncu_full.log
