"""Per source line of a kernel in an ncu report: instructions executed and stall samples, top lines.
python tools/ncu_source_lines.py report.ncu-rep [kernel-substring] [top]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; sub = sys.argv[2] if len(sys.argv) > 2 else ""; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur_file = cur_fn = None; hdr = None; acc = {}; order = []
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": cur_fn = r[1]; continue
    if r[0] == "Line No": hdr = r; iS = hdr.index("# Samples"); iI = hdr.index("Instructions Executed"); continue
    if hdr is None or not r[0] or sub not in (cur_fn or ""): continue
    if len(r) > len(hdr):  # the source text held a quote / comma the CSV writer did not escape: re-join it
        extra = len(r) - len(hdr)
        r = [r[0], ",".join(r[1:2 + extra])] + r[2 + extra:]
    key = (cur_fn[:60], cur_file, int(r[0]), r[1].strip()[:110])
    if key not in acc: acc[key] = [0, 0]; order.append(key)
    num = lambda x: int(x) if x not in ('-', '') else 0
    acc[key][0] += num(r[iI]); acc[key][1] += num(r[iS])
fns = {}
for k, v in acc.items():
    f = fns.setdefault(k[0], [0, 0]); f[0] += v[0]; f[1] += v[1]
for fn, (ti, ts) in fns.items():
    print("==", fn, "inst", ti, "samples", ts)
    ks = [k for k in order if k[0] == fn]
    ks.sort(key=lambda k: -acc[k][1])
    for k in ks[:top]:
        print("  %5.1f%% smp %5.1f%% inst  %s:%d  %s" % (100 * acc[k][1] / max(ts, 1), 100 * acc[k][0] / max(ti, 1), k[1], k[2], k[3]))
