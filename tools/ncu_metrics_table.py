"""Pivot an `ncu --csv --metrics ...` log into one row per kernel launch: python tools/ncu_metrics_table.py log.csv"""
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]; rows = rows[1:]
iid, ik, im, iv = hdr.index("ID"), hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
tab, order = {}, []
for r in rows:
    key = (int(r[iid]), r[ik].split("(")[0])
    if key not in tab: tab[key] = {}; order.append(key)
    tab[key][r[im]] = r[iv]
cols = ["gpu__time_duration.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__registers_per_thread"]
print("| id | kernel | time us | issue % | warps % | warp-inst M | L1 % | L2 % | dram rd MB | dram wr MB | grid | regs |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|")
tot = 0.0
for k in order:
    m = tab[k]
    f = lambda c: float(m.get(c, "0").replace(",", ""))
    t = f(cols[0]) / 1e3 if f(cols[0]) > 1e3 else f(cols[0])
    tot += t
    print("| %d | %s | %.1f | %.1f | %.1f | %.2f | %.1f | %.1f | %.2f | %.2f | %d | %d |" % (
        k[0], k[1], t, f(cols[1]), f(cols[2]), f(cols[3]) / 1e6, f(cols[4]), f(cols[5]), f(cols[6]) / 1e6 if f(cols[6]) > 1e4 else f(cols[6]),
        f(cols[7]) / 1e6 if f(cols[7]) > 1e4 else f(cols[7]), f(cols[8]), f(cols[9])))
print("\ntotal %.1f us" % tot)
