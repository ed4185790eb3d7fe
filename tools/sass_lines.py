#!/usr/bin/env python3
"""Static SASS size per source line/function for one kernel (nvdisasm -g -c output)."""
import re, sys, collections
txt = open(sys.argv[1]).read().split("\n")
kern = sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
inside = False; cur = ("?", 0); byline = collections.Counter(); byfile = collections.Counter(); total = 0
lab = None; bylabel = collections.Counter()
for l in txt:
    if l.startswith(".text."):
        inside = kern in l; continue
    if not inside: continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r'^(\$?[\w$.]+):', l)
    if m and not l.startswith(".L_x") and not l.startswith(".L"): lab = m.group(1)
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/', l):
        byline[cur] += 1; byfile[cur[0]] += 1; total += 1; bylabel[lab] += 1
print("kernel %s: %d instr = %.1f KB" % (kern, total, total * 16 / 1024))
for f, n in byfile.most_common(): print("  %-32s %6d %5.1f%%" % (f, n, 100 * n / total))
print("by function label:")
for f, n in bylabel.most_common(25): print("  %-70s %6d" % (str(f)[:70], n))
print("by line:")
for (f, ln), n in byline.most_common(top): print("  %-20s %5d %6d" % (f, ln, n))
