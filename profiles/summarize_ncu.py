"""Summarise an ncu report (run here, no GPU needed): key raw metrics + the hottest source lines.
    python profiles/summarize_ncu.py gpurun_out/prof_X.ncu-rep > profiles/X_summary.txt"""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__grid_size", "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__sass_average_branch_targets_threads_uniform.pct", "lts__t_sectors_srcunit_tex_op_read.sum", "l1tex__t_bytes.sum"]


def main(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print("== kernel:", d.get("Kernel Name"), " grid", d.get("Grid Size"), " block", d.get("Block Size"))
        for k in KEYS:
            if k in d:
                print(f"{k:95s} {d[k]:>16s} {units[hdr.index(k)]}")
    src = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    h = None
    insts = []
    for r in rows:
        if r and r[0] == "Address":
            h = r
            continue
        if h and len(r) == len(h):
            d = dict(zip(h, r))
            try:
                insts.append((int(d["Instructions Executed"]), float(d["Avg. Threads Executed"] or 0), int(d["# Samples"] or 0), d["Source"].strip()))
            except ValueError:
                pass
    total = sum(i[0] for i in insts) or 1
    tot_thr = sum(i[0] * i[1] for i in insts)
    print(f"\nSASS instructions: {len(insts)}  warp-insts executed: {total}  mean active threads/inst: {tot_thr / total:.2f}")
    by_op = {}
    for n, thr, smp, s in insts:
        op = s.split()[0] if not s.startswith("@") else s.split()[1]
        op = op.split(".")[0]
        a = by_op.setdefault(op, [0, 0.0, 0])
        a[0] += n; a[1] += n * thr; a[2] += smp
    print("\nopcode              warp-insts   share  avg-threads  stall-samples")
    for op, (n, t, smp) in sorted(by_op.items(), key=lambda kv: -kv[1][0])[:24]:
        print(f"{op:18s} {n:12d}  {100 * n / total:5.1f}%  {t / max(n, 1):6.2f}  {smp:8d}")


if __name__ == "__main__":
    main(sys.argv[1])
