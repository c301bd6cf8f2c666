"""Per-source-line summary of an ncu report: usage ncu_lines.py report.ncu-rep [min_pct]
(needs -lineinfo at compile time and --import-source on at capture time)."""
import csv, subprocess, sys, io
rep = sys.argv[1]; min_pct = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
cur_file, out, hdr = None, [], None
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if len(r) >= 2 and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr) or r[0] == "": continue
    d = dict(zip(hdr[4:], r[4:]))
    try:
        out.append((cur_file, int(r[0]), r[1].strip()[:90], int(d["Instructions Executed"] or 0),
                    int(d["Thread Instructions Executed"] or 0), int(d["# Samples"] or 0)))
    except ValueError:
        pass
ti = sum(o[3] for o in out) or 1; ts = sum(o[5] for o in out) or 1
print("total warp-inst %d, thread-inst %d (avg %.1f thr), samples %d" % (ti, sum(o[4] for o in out), sum(o[4] for o in out) / ti, ts))
for o in sorted(out, key=lambda o: -o[3]):
    if 100.0 * o[3] / ti < min_pct and 100.0 * o[5] / ts < min_pct: continue
    print("%5.1f%% inst %5.1f%% samp thr %4.1f  %s:%d  %s" % (100.0 * o[3] / ti, 100.0 * o[5] / ts, o[4] / max(o[3], 1), o[0], o[1], o[2]))
