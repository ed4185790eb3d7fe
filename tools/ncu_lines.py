#!/usr/bin/env python3
"""Per-source-line hot spots from `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv`."""
import csv, sys, os, collections
rows = csv.reader(open(sys.argv[1]))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None; hdr = None; out = []
for r in rows:
    if len(r) == 2 and r[0] == "File Path": cur = os.path.basename(r[1]); continue
    if r and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[2] == "-":
        d = {}
        for k, v in zip(hdr, r):
            d.setdefault(k, v)
        out.append((cur, int(r[0]), r[1], int(d["Instructions Executed"] or 0), int(d["# Samples"] or 0),
                    int(d["Thread Instructions Executed"] or 0)))
ti = sum(o[3] for o in out); ts = sum(o[4] for o in out)
print("total warp-instr %d, samples %d" % (ti, ts))
byfile = collections.Counter(); sf = collections.Counter()
for o in out: byfile[o[0]] += o[3]; sf[o[0]] += o[4]
for f, n in byfile.most_common(): print("%-18s instr %5.1f%%  samples %5.1f%%" % (f, 100.0 * n / ti, 100.0 * sf[f] / ts))
print()
for o in sorted(out, key=lambda o: -o[4])[:top]:
    print("%-16s %4d  smp %4.1f%%  ins %4.1f%%  thr/ins %4.1f | %s" % (o[0], o[1], 100.0 * o[4] / ts, 100.0 * o[3] / ti, o[5] / max(o[3], 1), o[2].strip()[:90]))
