#!/bin/bash
# GPU run M (1 x B200): the round's single-GPU evidence -- all GPU tests (verbose log), the default
# bench line, the reference arm, the ncu launch list and the `ncu --set full` captures of the three
# tile kernels.  ncu only after the same command has exited 0 without it.
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q -v > gpurun_out/r2_m_tests.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/r2_m_tests.log
timeout 600 python bench.py > gpurun_out/r2_m_bench.json 2> gpurun_out/r2_m_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r2_m_bench.json")); r=d["roofline"]
print(f"step {d['ms_per_step']:.4f} value {d['value']:.1f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} frac {r['frac']:.3f} side {r['side_kernels_ms']:.4f} e2e {d['e2e']['ms_per_step']:.4f}")
print("sustained", d.get("sustained")); print("cpu", d.get("cpu_baseline")); print("parity", d.get("parity")); print("clocks", d.get("clocks"))
PY
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_m_bench_reference.json 2>> gpurun_out/r2_m_bench.err
echo "reference rc=$?"; cat gpurun_out/r2_m_bench_reference.json
A="--steps 2 --warmup 3 --no-cpu-baseline --no-sustained --no-graph"
timeout 300 python bench.py $A > gpurun_out/r2_m_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 80 --csv \
    --log-file gpurun_out/r2_m_launches.csv python bench.py $A > gpurun_out/r2_m_ncu0.log 2>&1
echo "ncu launches rc=$?"
cap() { # name, kernel regex, extra env
  name=$1; k=$2; shift 2
  env "$@" timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f \
      -o gpurun_out/r2_m_$name python bench.py $A > gpurun_out/r2_m_ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
}
cap bwd tc_bwd_kernel A=1
cap fwd_sym tc_fwd_sym_kernel A=1
NTXENT_NO_SYM_FORWARD=1 timeout 300 python bench.py $A > gpurun_out/r2_m_plain2.log 2>&1 &&
cap fwd_full "tc_fwd_kernel" NTXENT_NO_SYM_FORWARD=1
ls -la gpurun_out/*.ncu-rep
