"""Instruction share by source-line range: ncu_ranges.py report.ncu-rep file:lo-hi:label ..."""
import csv, subprocess, sys, io
rep = sys.argv[1]
specs = []
for s in sys.argv[2:]:
    f, r, label = s.split(":"); lo, hi = r.split("-"); specs.append((f, int(lo), int(hi), label))
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
cur, hdr, agg, tot, tott = None, None, {}, 0, 0
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if len(r) >= 2 and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr) or r[0] == "": continue
    d = dict(zip(hdr[4:], r[4:]))
    try: ln, ie, te, sm = int(r[0]), int(d["Instructions Executed"] or 0), int(d["Thread Instructions Executed"] or 0), int(d["# Samples"] or 0)
    except ValueError: continue
    tot += ie; tott += te
    label = "other:%s" % cur
    for f, lo, hi, lb in specs:
        if cur == f and lo <= ln <= hi: label = lb; break
    a = agg.setdefault(label, [0, 0, 0]); a[0] += ie; a[1] += te; a[2] += sm
print("total warp-inst %.3e thread-inst %.3e" % (tot, tott))
for k, v in sorted(agg.items(), key=lambda x: -x[1][0]):
    print("%-28s %5.1f%% inst  thr %4.1f  samples %d" % (k, 100.0 * v[0] / tot, v[1] / max(v[0], 1), v[2]))
