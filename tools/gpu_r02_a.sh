#!/bin/bash
# round 2, call A: fixture of the oracle's w50 Fock matrix, GPU tests, whole-SCF parity, bench line, ncu full of the benched tensor kernels
mkdir -p gpurun_out
free -g | head -2; nproc; nvidia-smi topo -m | head -4
python tools/make_w50_fock_fixture.py > gpurun_out/a_fixture.log 2>&1; echo "fixture rc=$?"; tail -3 gpurun_out/a_fixture.log
python -m pytest tests -m "gpu and not slow" -q -rs -s > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/a_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/a_bench.json; tail -3 gpurun_out/a_bench.err
python tools/w50_scf_parity.py > gpurun_out/a_scf.log 2>&1; echo "scf parity rc=$?"; tail -c 1200 gpurun_out/a_scf.log
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/a_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:^k_exchange(_ws)?$|k_pp_half_transform' -s 6 -c 2 \
    -o gpurun_out/a_prof python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/a_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/a_ncu.log
