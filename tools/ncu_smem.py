"""Shared-memory wavefronts per CUDA source line from an `ncu --page source --print-source cuda,sass --csv` dump.
  python tools/ncu_smem.py src.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
cur, hdr, out = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or not r[0].isdigit():
        continue
    d = dict(zip(hdr[4:], r[4:]))

    def num(k):
        try:
            return float(d.get(k, "0"))
        except ValueError:
            return 0.0
    w, e = num("L1 Wavefronts Shared"), num("L1 Wavefronts Shared Excessive")
    if w:
        out.append((w, e, cur, int(r[0]), r[1].strip()[:90]))
tw = sum(o[0] for o in out) or 1
te = sum(o[1] for o in out)
print("shared wavefronts %d, excessive %d (%.1f%%)" % (tw, te, 100 * te / tw))
for w, e, f, n, src in sorted(out, key=lambda o: -o[0])[:top]:
    print("%5.1f%% wf  %5.1f%% of them excessive  %s:%d  %s" % (100 * w / tw, 100 * e / max(w, 1), f, n, src))
