"""Summarise an ncu report's SASS page: python tools/ncu_sass_summary.py report.ncu-rep [block]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; blk = int(sys.argv[2]) if len(sys.argv) > 2 else 100
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]; rows = rows[2:]
iS = hdr.index("# Samples") if "# Samples" in hdr else 2
iI = hdr.index("Instructions Executed")
tot = sum(int(r[iI]) for r in rows); samp = sum(int(r[iS]) for r in rows)
print("total inst", tot, "samples", samp, "n sass", len(rows))
for i in range(0, len(rows), blk):
    seg = rows[i:i + blk]
    ic = sum(int(r[iI]) for r in seg); sc = sum(int(r[iS]) for r in seg)
    ops = {}
    for r in seg:
        t = r[1].split(); op = t[1] if t[0].startswith('@') else t[0]
        ops[op] = ops.get(op, 0) + int(r[iI])
    top = sorted(ops.items(), key=lambda x: -x[1])[:5]
    print(i, "inst %.1f%%" % (100 * ic / tot), "samples %.1f%%" % (100 * sc / max(samp, 1)), [(k, round(100 * v / tot, 1)) for k, v in top])
