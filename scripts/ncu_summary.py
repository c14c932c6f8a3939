"""Turn an .ncu-rep (brought back in gpurun_out/) into the text summary committed under profiles/.

    python scripts/ncu_summary.py gpurun_out/x.ncu-rep profiles/x.md

Needs only the ncu CLI (no GPU).  Writes the headline counters per captured launch and, when the report was
taken with --import-source on from a -lineinfo build, the source lines ranked by stall samples / instructions.
"""
import csv
import io
import subprocess
import sys

RAW_KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__inst_executed.avg.per_cycle_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__warps_eligible.avg.per_cycle_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
]


def ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep, *args], capture_output=True, text=True).stdout


def main():
    rep, out = sys.argv[1], sys.argv[2]
    lines = [f"# ncu summary of `{rep.split('/')[-1]}`", ""]
    rows = list(csv.reader(io.StringIO(ncu(rep, "--page", "raw", "--csv"))))
    h, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[h.index("Kernel Name")]
        lines += [f"## {name[:110]}", "", f"grid {r[h.index('Grid Size')]} block {r[h.index('Block Size')]}", "",
                  "| metric | value | unit |", "|---|---|---|"]
        for k in RAW_KEYS:
            if k in h:
                i = h.index(k)
                lines.append(f"| {k} | {r[i]} | {units[i]} |")
        lines.append("")
    src = list(csv.reader(io.StringIO(ncu(rep, "--page", "source", "--print-source", "cuda,sass", "--csv"))))
    cur, agg, ts, ti = None, [], 0, 0
    for r in src:
        if len(r) >= 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if len(r) < 8 or r[0] in ("Line No", ""):
            continue
        try:
            ln, s, i = int(r[0]), int(r[4]), int(r[7])
        except ValueError:
            continue
        agg.append((cur, ln, r[1].strip()[:100], s, i))
        ts += s
        ti += i
    if agg and ts:
        lines += ["## source lines by warp-stall samples (last captured launch; inlined frames repeat their samples)", "",
                  "| file:line | samples % | inst % | source |", "|---|---|---|---|"]
        for a in sorted(agg, key=lambda t: -t[3])[:30]:
            lines.append(f"| {a[0]}:{a[1]} | {100 * a[3] / ts:.1f} | {100 * a[4] / max(ti, 1):.1f} | `{a[2]}` |")
        lines += ["", "## source lines by instructions executed", "", "| file:line | inst % | samples % | source |", "|---|---|---|---|"]
        for a in sorted(agg, key=lambda t: -t[4])[:20]:
            lines.append(f"| {a[0]}:{a[1]} | {100 * a[4] / max(ti, 1):.1f} | {100 * a[3] / ts:.1f} | `{a[2]}` |")
    open(out, "w").write("\n".join(lines) + "\n")
    print("wrote", out)


if __name__ == "__main__":
    main()
