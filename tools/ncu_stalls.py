#!/usr/bin/env python
"""Developer tool: headline counters and the per-issue stall breakdown of the first kernel in an .ncu-rep (no GPU needed)."""
import csv, io, subprocess, sys
for path in sys.argv[1:]:
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, u = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(h, r)); un = dict(zip(h, u))
        f = lambda k: float(d[k].replace(",", "")) if d.get(k) not in (None, "") else float("nan")
        print(f"== {path}: {d['Kernel Name'][:60]}")
        print(f"   time {f('gpu__time_duration.sum'):.1f} {un['gpu__time_duration.sum']}  DRAM rd {f('dram__bytes_read.sum'):.3f} {un['dram__bytes_read.sum']} wr {f('dram__bytes_write.sum'):.3f} {un['dram__bytes_write.sum']}"
              f"  regs {d['launch__registers_per_thread']}  issue active {f('smsp__issue_active.avg.pct_of_peak_sustained_active'):.1f}%  fp64 pipe {f('sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active'):.1f}%"
              f"  icc hit {f('sm__icc_request_hit_rate.pct'):.1f}%  L2 hit {f('lts__t_sector_hit_rate.pct'):.1f}%  inst {f('smsp__inst_executed.sum'):.4g}")
        st = [(k, f(k)) for k in h if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and "not_issued" not in k]
        print("   stalls per issue: " + ", ".join(f"{k.split('issue_stalled_')[1].split('_per_issue')[0]} {v:.2f}" for k, v in sorted(st, key=lambda x: -x[1])[:9]))
