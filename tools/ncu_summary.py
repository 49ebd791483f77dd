"""Turn ncu outputs into the small text summaries committed under profiles/.

    python tools/ncu_summary.py launches <launches.csv> [out.txt]    # --metrics gpu__time_duration.sum pass
    python tools/ncu_summary.py full <report.ncu-rep> [out.txt]      # --set full capture of one kernel
"""
import collections
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_active.avg.per_cycle_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fp64.sum", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_lsu.sum",
    "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum", "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
]


def launches(path, out):
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    agg = collections.OrderedDict()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        k = r["Kernel Name"].split("(")[0].replace("void ", "")
        v = float(r["Metric Value"].replace(",", ""))
        if r.get("Metric Unit") == "us":
            v *= 1e3
        elif r.get("Metric Unit") == "ms":
            v *= 1e6
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print("# ncu --metrics gpu__time_duration.sum --clock-control none launch list, aggregated per kernel", file=out)
    print("# (cold-cache, serialised launches: the SHARE is meaningful, not the absolute time)", file=out)
    print("%-64s %6s %12s %10s %7s" % ("kernel", "n", "total_us", "avg_us", "share"), file=out)
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-64s %6d %12.1f %10.1f %6.1f%%" % (k[:64], n, t / 1e3, t / n / 1e3, 100 * t / tot), file=out)
    print("%-64s %6d %12.1f" % ("TOTAL", sum(a[0] for a in agg.values()), tot / 1e3), file=out)


def full(path, out):
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    for vals in rows[2:]:
        d = dict(zip(hdr, vals))
        u = dict(zip(hdr, units))
        print("# ncu --set full --clock-control none: %s  grid %s block %s" % (
            d.get("Kernel Name", "?")[:90], d.get("Grid Size", "?"), d.get("Block Size", "?")), file=out)
        for k in KEYS:
            if k in d:
                print("%-82s %-10s %s" % (k, u[k], d[k]), file=out)
        try:
            rd = float(d["dram__bytes_read.sum"].replace(",", ""))
            wr = float(d["dram__bytes_write.sum"].replace(",", ""))
            mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            print("%-82s %-10s %.0f" % ("traffic = dram read + write", "byte",
                                        rd * mult[u["dram__bytes_read.sum"]] + wr * mult[u["dram__bytes_write.sum"]]), file=out)
        except Exception:
            pass
        print(file=out)


if __name__ == "__main__":
    mode, path = sys.argv[1], sys.argv[2]
    out = open(sys.argv[3], "w") if len(sys.argv) > 3 else sys.stdout
    (launches if mode == "launches" else full)(path, out)
