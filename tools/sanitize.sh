#!/bin/bash
# compute-sanitizer over the small parity cases (C1-sized and a C2-like size, both kernel families,
# index-vector positives, cross-view form).  ONE tool per call, as the profiling recipe asks:
#   bash tools/sanitize.sh memcheck   |   bash tools/sanitize.sh racecheck
# Writes gpurun_out/sanitizer_<tool>.log; the summary lines go to profiles/r2_sanitizer_<tool>.log.
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 900 compute-sanitizer --tool $TOOL --print-limit 20 python - > gpurun_out/sanitizer_$TOOL.log 2>&1 <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import ntxent_b200
from ntxent_b200 import capi
from conftest import seeded_z
for M, D, dtype, flags in [(512, 128, torch.float32, 0), (512, 128, torch.float32, capi.FLAG_FORCE_SIMT),
                           (1024, 128, torch.bfloat16, 0), (4096, 128, torch.bfloat16, 0), (1280, 64, torch.float32, 0),
                           (1024, 128, torch.bfloat16, capi.FLAG_CROSS_VIEW)]:
    z = torch.from_numpy(seeded_z(M, D)).cuda().to(dtype)
    loss, lse, dz = capi.forward_backward(z, 0.1 if dtype == torch.bfloat16 else 0.5, None, False, flags)
    torch.cuda.synchronize()
    print("ok", M, D, dtype, flags, float(loss))
z = torch.from_numpy(seeded_z(1024, 128)).cuda().to(torch.bfloat16)
lists = capi.inverse_positives((torch.arange(1024, device="cuda") + 3) % 1024)
print("pos", float(capi.forward_backward_pos(z, lists, 0.1)[0]))
(comm,) = capi.Communicator.init_all([0], 1024, 128, torch.bfloat16)
for ex, mode in ((capi.EXCHANGE_PUSH, capi.BWD_SYMMETRIC), (capi.EXCHANGE_NCCL, capi.BWD_CONTRACT)):
    print("comm", float(comm.forward_backward(z, 0.1, None, ex, mode)[0]))
torch.cuda.synchronize()
comm.destroy()
PY
echo "rc=$?" >> gpurun_out/sanitizer_$TOOL.log
grep -E "^ok|^pos|^comm|ERROR SUMMARY|RACECHECK SUMMARY|rc=|Error|error" gpurun_out/sanitizer_$TOOL.log | head -40
