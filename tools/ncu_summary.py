#!/usr/bin/env python
"""Summarise the ncu outputs of tools/gpu_profile.sh <tag> (run in the dev container, ncu -i needs no GPU):
   python tools/ncu_summary.py r1h > profiles/r1h_ncu_summary.txt"""
import collections, csv, io, os, subprocess, sys
tag = sys.argv[1]
out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
rows = [r for r in csv.reader(l for l in open(os.path.join(out, f"{tag}_launches.csv")) if l.startswith('"'))]
hdr, rows = rows[0], rows[1:]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows:
    tot[r[ki]] += float(r[vi].replace(",", "")) / 1e3
    cnt[r[ki]] += 1
total = sum(tot.values())
print(f"{tag}: ncu launch list of `python bench.py --steps 2 --warmup 1 --max-it 20 --no-e2e --no-cpu-baseline --no-parity` (8192^2, B200)")
print("ncu --metrics gpu__time_duration.sum --clock-control none -c 400   (cold-cache, serialised: compare SHARES)")
print(f"{len(rows)} launches, {total:.1f} us in kernels")
for k, v in tot.most_common():
    print(f"{k[:70]:70s} n={cnt[k]:4d} total={v:10.1f} us avg={v / cnt[k]:8.1f} share={v / total:.3f}")
print()
for rep in (f"{tag}_sor_persist", f"{tag}_momentum"):
    path = os.path.join(out, rep + ".ncu-rep")
    if not os.path.exists(path):
        continue
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(raw)))
    h, units = rr[0], rr[1]
    print(f"ncu --set full --clock-control none ({rep}.ncu-rep), per launch:")
    for r in rr[2:]:
        d = dict(zip(h, r))
        u = dict(zip(h, units))
        f = lambda k: float(d[k].replace(",", "")) if d.get(k) not in (None, "") else float("nan")
        dur = f("gpu__time_duration.sum") * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u["gpu__time_duration.sum"], 1.0)
        scale = {"byte": 1e-9, "Kbyte": 1e-6, "Mbyte": 1e-3, "Gbyte": 1.0, "Tbyte": 1e3}
        rd = f("dram__bytes_read.sum") * scale.get(u["dram__bytes_read.sum"], 1.0)
        wr = f("dram__bytes_write.sum") * scale.get(u["dram__bytes_write.sum"], 1.0)
        print(f"{d['Kernel Name'][:44]:44s} {dur:9.1f} us  DRAM read {rd:7.3f} GB write {wr:7.3f} GB  "
              f"DRAM {f('dram__cycles_active.avg.pct_of_peak_sustained_elapsed'):5.1f}% DRAM cycles active  issue active {f('smsp__issue_active.avg.pct_of_peak_sustained_active'):5.1f}%  "
              f"icc hit {f('sm__icc_request_hit_rate.pct'):5.1f}%  no_instruction/issue {f('smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio'):4.2f}  "
              f"long_scoreboard/issue {f('smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio'):4.2f}  regs {d.get('launch__registers_per_thread')}")
    print()
