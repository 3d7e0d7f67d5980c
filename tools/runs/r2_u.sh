#!/bin/bash
# GPU run U (1 x B200): ncu --set full of the two side kernels that take longest in situ.
mkdir -p gpurun_out
A="--steps 2 --warmup 3 --no-cpu-baseline --no-sustained --no-graph"
timeout 300 python bench.py $A > gpurun_out/r2_u_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
for k in bwd_epilogue_kernel fwd_sym_finalize_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f \
      -o gpurun_out/r2_u_$k python bench.py $A > gpurun_out/r2_u_ncu_$k.log 2>&1
  echo "ncu $k rc=$?"
done
