"""Summarise an .ncu-rep (read here, no GPU): key launch metrics + hottest source lines.
  python scripts/ncu_summary.py gpurun_out/x.ncu-rep [out.txt]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keys = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.avg.per_cycle_elapsed",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "l1tex__t_bytes.sum", "lts__t_bytes.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"]
out = []
for r in rows[2:]:
    out.append("== launch %s" % r[0])
    for k in keys:
        if k in hdr:
            i = hdr.index(k)
            out.append("%-86s %-12s %s" % (k, units[i], r[i][:110]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
# cuda,sass view: blocks per source file; rows are either a source line ("Line No","Source") followed by its SASS rows
agg = {}   # (file, line) -> [samples, warp_inst, thread_inst, text]
cur_file, cur = None, None
hdr2 = None
for r in csv.reader(io.StringIO(src)):
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr2 = r; continue
    if hdr2 is None:
        continue
    if r[0].strip():  # a source line row
        cur = (cur_file, int(r[0]))
        agg.setdefault(cur, [0, 0, 0, r[1].strip()[:105]])
        continue
    if cur is None:
        continue
    d = dict(zip(hdr2, r))
    try:
        agg[cur][0] += int(d.get("# Samples") or 0)
        agg[cur][1] += int(d.get("Instructions Executed") or 0)
        agg[cur][2] += int(d.get("Thread Instructions Executed") or 0)
    except ValueError:
        pass
tot = sum(v[0] for v in agg.values()) or 1
toti = sum(v[1] for v in agg.values()) or 1
out.append("== hottest source lines (total %d warp-stall samples, %d warp instructions); columns: %%samples, %%warp-inst, lanes/inst, file:line, source" % (tot, toti))
for (f, n), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:50]:
    out.append("%5.1f%% %5.1f%% %5.1f  %s:%d  %s" % (100.0 * v[0] / tot, 100.0 * v[1] / toti, (v[2] / v[1]) if v[1] else 0, f, n, v[3]))
txt = "\n".join(out)
print(txt)
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(txt + "\n")
