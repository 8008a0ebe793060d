"""Summarise an .ncu-rep: per-kernel headline metrics + hot-loop instruction mix (needs -lineinfo / --import-source)."""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "lts__t_sectors_op_write.sum", "lts__t_bytes.sum",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_drain_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio"]
for r in rows[2:]:
    print("-----")
    for w in want:
        if w in hdr:
            i = hdr.index(w)
            print(f"{w} = {r[i]} {units[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
kern, data, h = None, {}, None
for r in rows:
    if r and r[0] == "Kernel Name":
        kern = r[1]
        data[kern] = []
        continue
    if r and r[0] == "Address":
        h = r
        continue
    if kern and r:
        data[kern].append(r)
for k, v in data.items():
    iE, iS = h.index("Instructions Executed"), h.index("Source")
    iSt = h.index("Warp Stall Sampling (All Samples)")
    tot = sum(int(r[iE]) for r in v)
    mx = max(int(r[iE]) for r in v)
    body = [r for r in v if int(r[iE]) > 0.7 * mx]
    ops = collections.Counter()
    for r in body:
        t = r[iS].split()
        op = t[1] if t[0].startswith("@") else t[0]
        ops[op.split(".")[0]] += 1
    print("=====", k)
    print("total warp-instructions", tot, "hot-loop SASS instructions", len(body), "max per-instruction executions", mx)
    print("hot-loop opcode mix:", ops.most_common(30))
    top = sorted(v, key=lambda r: -int(r[iSt] or 0))[:14]
    print("top stall-sample instructions:")
    for r in top:
        print("   ", r[iSt], r[iS][:90])
