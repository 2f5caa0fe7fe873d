"""Summarise gpurun_out/ ncu artefacts into profiles/ (tracked): python tools/make_profiles.py <round-tag>
  * profiles/<tag>_bench_launches.csv / .md : the gpu__time_duration launch list of `bench.py` (per-kernel totals + shares)
  * profiles/<tag>_kernels_full.md          : one table row per `ncu --set full` capture (prof_*_<tag>f.ncu-rep)
  * profiles/<tag>_traffic.json             : DRAM bytes per launch of each captured kernel (bench.py roofline.traffic)
"""
import csv, glob, io, json, os, subprocess, sys
tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
os.makedirs(P, exist_ok=True)

src = os.path.join(G, "%s_bench_launches.csv" % tag)
if os.path.exists(src):
    rows = [r for r in csv.reader(open(src)) if len(r) > 10]
    hdr, rows = rows[0], rows[1:]
    ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot, cnt = {}, {}
    for r in rows:
        k = r[ik].split("(")[0]
        v = float(r[iv].replace(",", ""))
        v = v / 1e3 if r[iu] in ("ns", "nsecond") else (v * 1e3 if r[iu] in ("ms", "msecond") else v)
        tot[k] = tot.get(k, 0.0) + v
        cnt[k] = cnt.get(k, 0) + 1
    with open(os.path.join(P, "%s_bench_launches.csv" % tag), "w") as f:
        f.write(open(src).read())
    s = sum(tot.values())
    with open(os.path.join(P, "%s_bench_launches.md" % tag), "w") as f:
        f.write("# %s — ncu launch list of `python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-matchers --latency-frames 2`\n\n" % tag)
        f.write("`ncu --metrics gpu__time_duration.sum --clock-control none -c 400` (first 400 launches; serialised, cold-cache "
                "replays: compare SHARES with bench.py `stages_ms_per_launch`, not absolutes).  Raw list: `%s_bench_launches.csv`.\n\n" % tag)
        f.write("| kernel | launches | total us | share | mean us |\n|---|---|---|---|---|\n")
        for k in sorted(tot, key=lambda k: -tot[k]):
            f.write("| %s | %d | %.1f | %.1f %% | %.1f |\n" % (k, cnt[k], tot[k], 100 * tot[k] / s, tot[k] / cnt[k]))

src = os.path.join(G, "%s_step_metrics_ncu.csv" % tag)
if os.path.exists(src):
    with open(os.path.join(P, "%s_step_metrics_ncu.csv" % tag), "w") as f:
        f.write(open(src).read())
    tab = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_metrics_table.py"), src], capture_output=True, text=True).stdout
    with open(os.path.join(P, "%s_step_launches.md" % tag), "w") as f:
        f.write("# %s — per-launch ncu metrics of one device-resident bench step\n\n" % tag)
        f.write("Command: `ncu --metrics gpu__time_duration.sum,... --clock-control none -c 60 --csv python tools/one_step.py 64 1` "
                "(64 KITTI stereo pairs: left handle, right handle, stereo).  Times are serialised, cold-cache ncu replays: compare "
                "SHARES with `bench.py` `stages_ms_per_launch`, not absolutes.\n\n")
        f.write(tab)

want = ["gpu__time_duration.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "launch__grid_size", "launch__block_size", "launch__registers_per_thread"]
traffic = {}
lines = []
for rep in sorted(glob.glob(os.path.join(G, "prof_*_%sf.ncu-rep" % tag))):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    krows = [r for r in rows[2:] if len(r) == len(hdr)]  # one row per captured launch (FAST: its two launches)
    first = True
    tot = {"dram": 0.0, "time": 0.0, "winst": 0.0}
    for r in krows:
        name = r[hdr.index("Kernel Name")].split("(")[0]
        def val(k):
            if k not in hdr: return float("nan")
            v = float(r[hdr.index(k)].replace(",", "")); u = units[hdr.index(k)]
            scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "ns": 1e-3, "us": 1.0, "ms": 1e3}.get(u, 1.0)
            return v * scale
        v = {k: val(k) for k in want}
        tot["dram"] += v["dram__bytes_read.sum"] + v["dram__bytes_write.sum"]
        tot["time"] += v["gpu__time_duration.sum"]
        tot["winst"] += v["smsp__inst_executed.sum"]
        if first:
            busy = v["smsp__issue_active.avg.pct_of_peak_sustained_active"]
            first = False
        label = name if len(krows) == 1 else "%s (launch %d of %d)" % (name, krows.index(r) + 1, len(krows))
        lines.append("| %s | %.1f | %.1f | %.1f | %.1f | %.1f / %.1f / %.1f / %.1f | %.1f / %.1f / %.1f | %.1f + %.1f | %d x %d | %d |" % (
            label, v[want[0]], v[want[1]], v[want[2]], v[want[3]] / 1e6, v[want[4]], v[want[5]], v[want[6]], v[want[7]], v[want[8]],
            v[want[9]], v[want[10]], v[want[11]] / 1e6, v[want[12]] / 1e6, v[want[13]], v[want[14]], v[want[15]]))
    # per STAGE: the sum over the stage's captured launches (bench.py divides by the stage's live time)
    if len(krows) > 1:  # the launches of one stage are instantiations of one template: key without "void " / "<...>"
        name = name.replace("void ", "").split("<")[0]
    traffic[name] = {"dram_bytes_per_launch": tot["dram"], "images": 64, "time_us": tot["time"],
                     "warp_inst_per_launch": tot["winst"], "issue_slots_busy_pct": busy, "launches_captured": len(krows)}
if lines:
    with open(os.path.join(P, "%s_kernels_full.md" % tag), "w") as f:
        f.write("# %s — `ncu --set full --clock-control none --import-source on` of each kernel, one launch over 64 KITTI images\n\n" % tag)
        f.write("Command per kernel: `ncu --set full ... -k regex:<kernel> -c 1 python tools/one_step.py 64 1` (the first launch of the "
                "kernel in one device-resident bench step, both launches for FAST; the .ncu-rep files stay in gpurun_out/, scratch).\n\n")
        f.write("| kernel | time us | issue slots busy % | warps active % | warp-inst M | pipe ALU / FMA / LSU / XU % | L1 / L2 / DRAM % of peak | DRAM read + write MB | grid x block | regs |\n")
        f.write("|---|---|---|---|---|---|---|---|---|---|\n")
        f.write("\n".join(lines) + "\n")
    json.dump(traffic, open(os.path.join(P, "%s_traffic.json" % tag), "w"), indent=1)
print("profiles written:", sorted(os.listdir(P)))
