"""Per-kernel summary (markdown table) of an .ncu-rep captured with --set full: one row per captured launch with the counters the
roofline discussion needs. usage: python tools/ncu_summary.py report.ncu-rep [more reports ...] > profiles/SUMMARY_kernels.md"""
import csv
import io
import subprocess
import sys

COLS = [
    ("gpu__time_duration.sum", "ms", lambda v, u: "%.3f" % (v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1.0))),
    ("launch__registers_per_thread", "regs", lambda v, u: "%d" % v),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %", lambda v, u: "%.1f" % v),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %", lambda v, u: "%.1f" % v),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "lanes/inst", lambda v, u: "%.1f" % v),
    ("smsp__inst_executed.sum", "warp inst (M)", lambda v, u: "%.1f" % (v / 1e6)),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU pipe %", lambda v, u: "%.1f" % v),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "LSU pipe %", lambda v, u: "%.1f" % v),
    ("dram__bytes_read.sum", "DRAM rd GB", lambda v, u: "%.3f" % (v * {"byte": 1e-9, "Kbyte": 1e-6, "Mbyte": 1e-3, "Gbyte": 1.0, "Tbyte": 1e3}.get(u, 1e-9))),
    ("dram__bytes_write.sum", "DRAM wr GB", lambda v, u: "%.3f" % (v * {"byte": 1e-9, "Kbyte": 1e-6, "Mbyte": 1e-3, "Gbyte": 1.0, "Tbyte": 1e3}.get(u, 1e-9))),
    ("dram__bytes.sum.per_second", "DRAM GB/s", lambda v, u: "%.0f" % (v * {"byte/s": 1e-9, "Kbyte/s": 1e-6, "Mbyte/s": 1e-3, "Gbyte/s": 1.0, "Tbyte/s": 1e3}.get(u, 1.0))),
    ("lts__t_sectors.sum", "L2 sectors (M)", lambda v, u: "%.1f" % (v / 1e6)),
    ("lts__t_sector_hit_rate.pct", "L2 hit %", lambda v, u: "%.1f" % v),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit %", lambda v, u: "%.1f" % v),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_sb", lambda v, u: "%.1f" % v),
    ("smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "stall no_inst", lambda v, u: "%.1f" % v),
    ("smsp__average_warps_active_per_inst_executed.ratio", "cycles/inst/warp", lambda v, u: "%.1f" % v),
]


def rows_of(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    if len(rows) < 3:
        return
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = {}
        for k, u, v in zip(hdr, units, r):
            d[k] = (v, u)
        yield d


def main():
    print("| kernel (launch) | " + " | ".join(c[1] for c in COLS) + " | L2 GB/s |")
    print("|---|" + "---|" * (len(COLS) + 1))
    for rep in sys.argv[1:]:
        for d in rows_of(rep):
            name = d.get("Kernel Name", ("?", ""))[0]
            name = name.split("(")[0].replace("void ", "")
            cells = []
            for key, _, fmt in COLS:
                v, u = d.get(key, ("", ""))
                try:
                    cells.append(fmt(float(v.replace(",", "")), u))
                except ValueError:
                    cells.append("-")
            try:   # L2 throughput: sectors x 32 B over the kernel's duration
                sec = float(d["lts__t_sectors.sum"][0].replace(",", ""))
                t, tu = d["gpu__time_duration.sum"]
                ms = float(t.replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(tu, 1.0)
                cells.append("%.0f" % (sec * 32 / (ms * 1e-3) / 1e9))
            except (KeyError, ValueError, ZeroDivisionError):
                cells.append("-")
            print("| `%s` (%s) | " % (name[:60], d.get("ID", ("", ""))[0]) + " | ".join(cells) + " |")


if __name__ == "__main__":
    main()
