"""Summarise ncu --set full reports (every kernel launch in them) into markdown.

usage: ncu_summary.py out.md title rep1.ncu-rep [rep2 ...]
Also prints, per report, 'kernel<TAB>dram_bytes' for profiles/dram_traffic.json."""
import csv
import io
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
]
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}


def main():
    out, title, reps = sys.argv[1], sys.argv[2], sys.argv[3:]
    md = ["# %s" % title, "",
          "per launch; times under ncu are cold-cache and serialised (compare "
          "shares, not absolutes)", ""]
    for rep in reps:
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"],
                             capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(txt)))
        if len(rows) < 3:
            continue
        hdr, units = rows[0], rows[1]
        for vals in rows[2:]:
            name = vals[hdr.index("Kernel Name")]
            md.append("## %s" % name[:110])
            dram = 0.0
            for k in KEEP:
                if k in hdr:
                    j = hdr.index(k)
                    md.append("- %s = %s %s" % (k, vals[j], units[j]))
                    if k.startswith("dram__bytes"):
                        dram += float(vals[j]) * SCALE.get(units[j], 1.0)
            md.append("- dram bytes read+write = %.4g GB" % (dram / 1e9))
            md.append("")
            print("%s\t%.0f" % (name.split("(")[0], dram))
    with open(out, "w") as f:
        f.write("\n".join(md))


if __name__ == "__main__":
    main()
