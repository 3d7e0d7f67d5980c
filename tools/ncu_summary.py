"""Turns `ncu --set full` reports (gpurun_out/*.ncu-rep) into the short CSV summaries committed
under profiles/: the metrics DESIGN.md and bench.py quote (duration, DRAM bytes, pipe activity, issue
slots, registers, stall picture) and the dynamic instruction mix from the source page.

    python tools/ncu_summary.py gpurun_out/r2_x_bwd.ncu-rep profiles/r2_ncu_bwd_summary.csv
"""
import collections
import csv
import io
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.max", "sm__cycles_elapsed.max.per_second",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.avg.per_cycle_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
]


def ncu(args):
    return subprocess.run(["ncu", *args], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout


def main(rep, out):
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units, vals = rows[0], rows[1], rows[2]
    lines = [("kernel", "", vals[hdr.index("Kernel Name")][:120])]
    for k in WANT:
        if k in hdr:
            i = hdr.index(k)
            lines.append((k, units[i], vals[i]))
    for i, h in enumerate(hdr):  # the stall picture: average warps per issue stalled on <reason>
        if h.startswith("smsp__average_warp") and "issue_stalled" in h and h.endswith("_per_issue_active.ratio"):
            lines.append((h, units[i], vals[i]))
    src = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "source", "--csv"]))))
    if len(src) > 2:
        h = src[1]
        ai, ei = h.index("Source"), h.index("Instructions Executed")
        mix, total = collections.Counter(), 0
        for r in src[2:]:
            if len(r) <= ei:
                continue
            t = r[ai].split()
            if not t:
                continue
            op = (t[1] if t[0].startswith("@") and len(t) > 1 else t[0]).split(".")[0]
            e = int(r[ei] or 0)
            mix[op] += e
            total += e
        lines.append(("instructions_executed_total (warp level)", "inst", str(total)))
        for op, e in mix.most_common(14):
            lines.append((f"inst_mix.{op}", "%", f"{100.0 * e / max(total, 1):.1f}"))
    with open(out, "w") as f:
        for k, u, v in lines:
            f.write(f"{k},{u},{v}\n")
    print(f"{out}: {len(lines)} lines")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
