"""tools/sass_calls.py <nvdisasm -g -c output> <kernel substring> <ncu --page source --print-source sass --csv> [min per warp-iteration]
Lists the CALL / RET instructions of a kernel that actually execute (out-of-line slow paths of division, square root, reciprocal
and libm), with their source line and executed count. A slow path that runs on every event is a bug in the calling code."""
import csv
import re
import sys

sass, kern, ncu = sys.argv[1:4]
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
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ie = h.index("Instructions Executed")
base, tot = None, 0
hits = []
for r in rows[hi + 1:]:
    if not r or not r[0].startswith("0x"):
        if base is not None:
            break
        continue
    a = int(r[0], 16)
    if base is None:
        base = a
    src, op = line_of.get(a - base, ((None, 0), "?"))
    e = int(r[ie] or 0)
    tot += e
    if ("CALL" in op or "RET" in op) and e > 0:
        hits.append((e, src, op[:100]))
print("total warp instructions", tot)
for e, src, op in sorted(hits, key=lambda t: -t[0])[:40]:
    print("%12d  %5.2f%%  %s:%s  %s" % (e, 100.0 * e / tot, src[0] if src else "?", src[1] if src else 0, op))
