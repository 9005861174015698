"""per-kernel summary of an ncu report (--page raw --csv): duration, DRAM / L2 throughput, issue
slots, occupancy, registers — the columns the profile summaries quote"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
want = [("gpu__time_duration.sum", "dur"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("dram__bytes.sum.per_second", "dram_B/s"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("lts__t_sectors.sum", "l2_MB"), 
        ("lts__t_sector_hit_rate.pct", "l2hit%"),
        ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "st_long_sb"),
        ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "st_barrier"),
        ("smsp__average_warps_issue_stalled_membar_per_issue_active.ratio", "st_membar"),
        ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "st_wait"),
        ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "st_short_sb"),
        ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "st_lg_thr"),
        ("smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "st_branch")]
print("| kernel | " + " | ".join(n for _, n in want) + " |")
print("|---|" + "---|" * len(want))
for r in data:
    name = re.sub(r"\(.*", "", r[col["Kernel Name"]]).split("::")[-1]
    vals = []
    for m, n in want:
        if m in col:
            v, u = r[col[m]], units[col[m]]
            try:
                f = float(v.replace(",", ""))
                if n == "dur":
                    f = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6}.get(u, 1.0) * f
                    vals.append(f"{f:.1f} us")
                elif n in ("dram_rd", "dram_wr", "l2_bytes"):
                    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
                    vals.append(f"{f * mult / 1e6:.2f} MB")
                elif n in ("l2_B/s", "dram_B/s"):
                    mult = {"byte/s": 1, "Kbyte/s": 1e3, "Mbyte/s": 1e6, "Gbyte/s": 1e9, "Tbyte/s": 1e12}.get(u, 1)
                    vals.append(f"{f * mult / 1e9:.0f} GB/s")
                elif n == "l2_MB":
                    vals.append(f"{f * 32 / 1e6:.1f} MB")
                else:
                    vals.append(f"{f:.4g}")
            except ValueError:
                vals.append(v)
        else:
            vals.append("-")
    print(f"| {name} | " + " | ".join(vals) + " |")
