#!/bin/bash
# GPU run B (1 x B200): GPU tests with the upper-triangle forward, bench A/B lines, drop-tile proof.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_b_tests.log 2>&1; echo "tests rc=$?"
tail -25 gpurun_out/r2_b_tests.log
B="--no-cpu-baseline --no-sustained"
timeout 300 python bench.py $B > gpurun_out/r2_b_bench_sym.json 2> gpurun_out/r2_b_bench.err; echo "bench rc=$?"
cat gpurun_out/r2_b_bench_sym.json; tail -3 gpurun_out/r2_b_bench.err
NTXENT_NO_SYM_FORWARD=1 timeout 300 python bench.py $B > gpurun_out/r2_b_bench_nosym.json 2>> gpurun_out/r2_b_bench.err
cat gpurun_out/r2_b_bench_nosym.json
NTXENT_NO_SYM_FORWARD=1 NTXENT_NO_PDL=1 timeout 300 python bench.py $B > gpurun_out/r2_b_bench_nosym_nopdl.json 2>> gpurun_out/r2_b_bench.err
cat gpurun_out/r2_b_bench_nosym_nopdl.json
timeout 300 python tools/drop_tile_proof.py > gpurun_out/r2_b_droptile.log 2>&1; cat gpurun_out/r2_b_droptile.log
