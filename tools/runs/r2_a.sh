#!/bin/bash
# GPU run A (1 x B200): GPU tests, default bench line, the drop-tile proof, the ncu launch list.
mkdir -p gpurun_out
nproc > gpurun_out/r2_a_env.log; nvidia-smi -L >> gpurun_out/r2_a_env.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_a_tests.log 2>&1; echo "tests rc=$?"
tail -15 gpurun_out/r2_a_tests.log
timeout 400 python bench.py > gpurun_out/r2_a_bench.json 2> gpurun_out/r2_a_bench.err; echo "bench rc=$?"
cat gpurun_out/r2_a_bench.json; tail -3 gpurun_out/r2_a_bench.err
timeout 200 python tools/drop_tile_proof.py > gpurun_out/r2_a_droptile.log 2>&1; cat gpurun_out/r2_a_droptile.log
ARGS="--steps 3 --warmup 3 --no-cpu-baseline --no-sustained --no-graph"
timeout 300 python bench.py $ARGS > gpurun_out/r2_a_plain.log 2>&1 &&
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 60 --csv \
    --log-file gpurun_out/r2_a_launches.csv python bench.py $ARGS > gpurun_out/r2_a_ncu.log 2>&1
echo "ncu rc=$?"
