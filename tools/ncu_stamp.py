"""Write profiles/kernel_evidence.json from an ncu raw page: per hot kernel the DRAM traffic per launch, the issue-slot
and fp32-pipe utilisation, stamped with the hash of the kernel's sources and the commit, so that bench.py can tell
whether the evidence belongs to the library it is timing.

  ncu -i rep.ncu-rep --page raw --csv > raw.csv
  python tools/ncu_stamp.py raw.csv profiles/r02_ncu_bench.txt        (second argument: the summary the numbers are in)
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}

    def val(r, key, want_unit=None):
        v = float(r[idx[key]].replace(",", ""))
        u = units[idx[key]]
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(u, 1.0)
        return v * scale

    commit = subprocess.check_output(["git", "rev-parse", "--short", "HEAD"], cwd=ROOT, text=True).strip()
    acc = {}
    for r in rows[2:]:
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "").split("<")[0].strip()
        if name not in bench.KERNEL_SOURCES:
            continue
        a = acc.setdefault(name, dict(n=0, traffic=0.0, issue=0.0, fma=0.0))
        a["n"] += 1
        a["traffic"] += val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")
        a["issue"] += val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active")
        a["fma"] += val(r, "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active")
    out = dict(_how="ncu --set full --clock-control none; mean over the captured launches of each kernel; "
                    "traffic = dram__bytes_read.sum + dram__bytes_write.sum per launch", kernels={})
    for name, a in acc.items():
        out["kernels"][name] = dict(traffic=a["traffic"] / a["n"], issue_frac=a["issue"] / a["n"] / 100,
                                    fma_pipe_frac=a["fma"] / a["n"] / 100, launches=a["n"],
                                    sources_sha=bench.sources_sha(name), commit=commit,
                                    profile=sys.argv[2] if len(sys.argv) > 2 else None)
    json.dump(out, open(os.path.join(ROOT, "profiles", "kernel_evidence.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
