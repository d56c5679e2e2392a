"""Development aid: turn gpurun_out/launches.csv (ncu launch list) and gpurun_out/*.ncu-rep into the
text summaries committed under profiles/.  usage: summarise_profiles.py <tag>   (e.g. r01b)"""
import csv, collections, subprocess, sys, os, io
tag = sys.argv[1]
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = os.path.join(R, "profiles")
# --- launch list: per kernel count / total / share
rows = [r for r in csv.reader(open(os.path.join(R, "gpurun_out", "launches.csv"))) if r and not r[0].startswith("==")]
h = rows[0]; ik = h.index("Kernel Name"); iv = h.index("Metric Value"); iu = h.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    v = float(r[iv].replace(",", ""))
    v = v / 1000.0 if r[iu] in ("ns", "nsecond") else v          # -> us
    a = agg.setdefault(r[ik], [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
with open(os.path.join(out, tag + "_launches_summary.txt"), "w") as f:
    f.write("ncu --metrics gpu__time_duration.sum --clock-control none -c 400, python bench.py --steps 2 --warmup 3 --no-cpu-baseline\n")
    f.write("(cold-cache, serialised launches: shares matter, not absolutes)\n\n%-70s %6s %12s %7s\n" % ("kernel", "count", "total us", "share"))
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write("%-70s %6d %12.1f %6.1f%%\n" % (k[:70], n, t, 100 * t / tot))
# --- full captures: key metrics per kernel
KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.per_cycle_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__cycles_active.avg", "sm__cycles_elapsed.max"]
for rep in sys.argv[2:]:
    path = os.path.join(R, "gpurun_out", rep + ".ncu-rep")
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(raw)))
    hh = rr[0]
    with open(os.path.join(out, "%s_%s_summary.txt" % (tag, rep)), "w") as f:
        f.write("ncu --set full --clock-control none --import-source on (%s.ncu-rep), one block per captured launch\n" % rep)
        for v in rr[2:]:
            d = dict(zip(hh, v))
            f.write("\n== %s\n" % d.get("Kernel Name", ""))
            for k in KEYS:
                if k in d: f.write("  %-72s %s %s\n" % (k, d[k], rr[1][hh.index(k)]))
            st = [(n.replace("smsp__pcsamp_warps_issue_stalled_", ""), float(d[n])) for n in hh
                  if n.startswith("smsp__pcsamp_warps_issue_stalled_") and "not_issued" not in n and d[n] not in ("", "n/a")]
            tt = sum(x for _, x in st) or 1
            f.write("  stall samples: " + ", ".join("%s %.0f%%" % (n, 100 * x / tt) for n, x in sorted(st, key=lambda kv: -kv[1])[:7]) + "\n")
