"""tools/sass_by_line.py <nvdisasm -g -c output> <kernel mangled-name substring> <ncu --page source --print-source sass --csv>
Joins ncu's per-SASS-instruction counters with nvdisasm's line markers and prints executed instructions and stall samples
per source line (inlined functions are attributed to the innermost file:line nvdisasm reports)."""
import csv
import re
import sys
from collections import defaultdict

sass, kern, ncu = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
line_of = {}
cur, on = None, False
for ln in open(sass):
    if ln.startswith("//--------------------- .text."):
        on = kern in ln
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        line_of[int(m.group(1), 16)] = (cur, m.group(2).strip())
rows = list(csv.reader(open(ncu)))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
ia, ie, it_, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
base = None
per_line = defaultdict(lambda: [0, 0, 0])
per_file = defaultdict(lambda: [0, 0, 0])
tot = [0, 0, 0]
for r in rows[hdr_i + 1:]:
    if not r or not r[0].startswith("0x"):
        if base is not None:
            break  # first launch only
        continue
    a = int(r[0], 16)
    if base is None:
        base = a
    src = line_of.get(a - base, ((None, 0), "?"))[0] or ("?", 0)
    v = [int(r[ie] or 0), int(r[it_] or 0), int(r[isamp] or 0)]
    for k in range(3):
        per_line[src][k] += v[k]
        per_file[src[0]][k] += v[k]
        tot[k] += v[k]
print("total warp-inst %d thread-inst %d samples %d" % tuple(tot))
print("-- by file")
for f, v in sorted(per_file.items(), key=lambda kv: -kv[1][0]):
    print("%-22s inst %6.2f%%  samples %6.2f%%" % (f, 100.0 * v[0] / tot[0], 100.0 * v[2] / max(tot[2], 1)))
print("-- by line (top %d by executed instructions)" % top)
for s, v in sorted(per_line.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-22s:%-5d inst %6.2f%%  samples %6.2f%%" % (s[0], s[1], 100.0 * v[0] / tot[0], 100.0 * v[2] / max(tot[2], 1)))
