#!/usr/bin/env python
"""Summarises an .ncu-rep (read here, no GPU needed) into profiles/<tag>_summary.md:
per captured launch the duration, DRAM bytes, throughput percentages, instruction counts and the top
stall reasons.  usage: scripts/ncu_summary.py gpurun_out/prof_X.ncu-rep profiles/X_summary.md [launches.csv]"""
import csv
import io
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__maximum_warps_per_active_cycle_pct",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_fp64.sum", "sm__inst_executed_pipe_lsu.sum",
    "lts__t_sectors.sum", "lts__t_sectors_op_atom.sum", "lts__t_sectors_op_red.sum",
    "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sector_hit_rate.pct",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
]


def raw_rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2:]


def main():
    rep, dst = sys.argv[1], sys.argv[2]
    hdr, units, rows = raw_rows(rep)
    lines = ["# ncu summary of `%s`" % rep.split("/")[-1], "",
             "Captured with `ncu --set full --clock-control none --import-source on` (scripts/prof1.sh); per-launch values.",
             "Numbers under a profiler are evidence of WHERE time and traffic go, not bench values.", ""]
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio")]
    if not stall_cols:
        stall_cols = [i for i, h in enumerate(hdr) if h.startswith("smsp__average_warp") and "stalled" in h and h.endswith(".ratio")]
    ki = hdr.index("Kernel Name")
    for r in rows:
        lines.append("## `%s`" % r[ki][:110])
        lines.append("")
        lines.append("| metric | value | unit |")
        lines.append("|---|---|---|")
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                lines.append("| %s | %s | %s |" % (w, r[i], units[i]))
        try:
            mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
            ri, wi = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
            rd = float(r[ri].replace(",", "")) * mult[units[ri]]
            wr = float(r[wi].replace(",", "")) * mult[units[wi]]
            lines.append("| **traffic (dram read + write)** | %.6g | Gbyte |" % ((rd + wr) / 1e9))
        except Exception:
            pass
        st = []
        for i in stall_cols:
            try:
                st.append((float(r[i].replace(",", "")), hdr[i]))
            except ValueError:
                pass
        st.sort(reverse=True)
        lines.append("")
        lines.append("Top warp stall reasons (warps stalled per issue-active cycle):")
        lines.append("")
        for v, n in st[:6]:
            short = n.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")
            lines.append("* %s: %.2f" % (short, v))
        lines.append("")
    if len(sys.argv) > 3:
        agg = {}
        total = 0.0
        with open(sys.argv[3]) as f:
            rd = [x for x in csv.reader(f) if len(x) > 5]
        h = rd[0]
        kn, mv = h.index("Kernel Name"), h.index("Metric Value")
        for x in rd[1:]:
            try:
                v = float(x[mv].replace(",", ""))
            except ValueError:
                continue
            name = x[kn].split("(")[0]
            a = agg.setdefault(name, [0, 0.0])
            a[0] += 1
            a[1] += v
            total += v
        lines.append("## launch list (`%s`: every launch of the same command, gpu__time_duration.sum)" % sys.argv[3].split("/")[-1])
        lines.append("")
        lines.append("| kernel | launches | total | share |")
        lines.append("|---|---|---|---|")
        for name, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            lines.append("| `%s` | %d | %.0f | %.1f %% |" % (name[:80], c, v, 100 * v / total))
        lines.append("")
    open(dst, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
