#!/bin/bash
# GPU run O (1 x B200): pipeline trace of the backward (two weight warps of one scheduler), with and
# without a start-up phase skew.
mkdir -p gpurun_out
for s in 0 400 800 100000; do
  echo "=== skew $s"
  NTXENT_BWD_SKEW=$s NTXENT_BWD_SKEW_PATTERN=0 timeout 200 python tools/trace_bwd.py > gpurun_out/r2_o_trace_$s.log 2>&1; echo "rc=$?"
  grep -v "^tile  [0-5]" gpurun_out/r2_o_trace_$s.log | tail -28
done
