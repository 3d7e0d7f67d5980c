#!/bin/bash
# GPU run V256 (1 x B200): ncu --set full of the D = 256 backward tile kernel (aliased-W layout, one S buffer).
mkdir -p gpurun_out
A="--M 65536 --D 256 --steps 2 --warmup 3 --no-cpu-baseline --no-sustained --no-graph"
timeout 300 python bench.py $A > gpurun_out/r2_v256_plain.json 2> gpurun_out/r2_v256.err || { echo "plain run failed"; tail -5 gpurun_out/r2_v256.err; exit 1; }
python -c "import json; d=json.load(open('gpurun_out/r2_v256_plain.json')); r=d['roofline']; print('step', d['ms_per_step'], 'bwd', r['ms_per_launch'], 'frac', r['frac'], 'fwd', r['fwd_kernel']['ms_per_launch'])"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_bwd_kernel -s 3 -c 1 -f \
    -o gpurun_out/r2_v256_bwd python bench.py $A > gpurun_out/r2_v256_ncu.log 2>&1
echo "ncu rc=$?"
