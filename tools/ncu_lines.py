"""Aggregate an `ncu --page source --print-source cuda,sass --csv` dump per CUDA source line.
  ncu -i rep.ncu-rep --page source --print-source cuda,sass --csv --kernel-name K > src.csv
  python tools/ncu_lines.py src.csv [top]
Prints the lines with the most stall samples and the most executed warp instructions."""
import csv
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(open(path)))
cur_file, hdr, lines = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or r[0] in ("Function Name", "") or not r[0].isdigit():
        continue
    d = dict(zip(hdr[4:], r[4:]))

    def num(k):
        try:
            return float(d.get(k, "0"))
        except ValueError:
            return 0.0
    stalls = {k: num(k) for k in d if k.startswith("stall_") and "Not Issued" not in k}
    lines.append((cur_file, int(r[0]), r[1].strip()[:90], num("# Samples"), num("Instructions Executed"), stalls))
tot_s = sum(l[3] for l in lines) or 1
tot_i = sum(l[4] for l in lines) or 1
print("total samples %d, warp instructions %d" % (tot_s, tot_i))
print("---- by samples")
for f, n, src, s, i, st in sorted(lines, key=lambda l: -l[3])[:top]:
    main = sorted(st.items(), key=lambda kv: -kv[1])[:2]
    print("%5.1f%% %5.1f%%i %s:%d  %s   [%s]" % (100 * s / tot_s, 100 * i / tot_i, f, n, src,
                                                ", ".join("%s %d" % (k[6:], v) for k, v in main)))
print("---- by instructions")
for f, n, src, s, i, st in sorted(lines, key=lambda l: -l[4])[:top]:
    print("%5.1f%%i %5.1f%% %s:%d  %s" % (100 * i / tot_i, 100 * s / tot_s, f, n, src))
