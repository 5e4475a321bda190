"""Top stall sites of a kernel from an .ncu-rep (SASS view): python tools/ncu_hot.py <file.ncu-rep> [topN]"""
import csv, io, subprocess, sys
txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(io.StringIO(txt)))
hdr = None; data = []; kern = 0
for r in rows:
    if r and r[0] == "Kernel Name":
        kern += 1
        continue
    if r and r[0] == "Address":
        hdr = r; continue
    if hdr and kern == 1 and len(r) >= 5:
        data.append(r)
i_src, i_s, i_ex = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
tot = sum(int(r[i_s] or 0) for r in data)
print("total samples", tot, "instructions", len(data))
for idx, r in sorted(enumerate(data), key=lambda t: -int(t[1][i_s] or 0))[:n]:
    print(f"{idx:5d} {int(r[i_s]):7d} {100*int(r[i_s])/tot:5.1f}%  exec {r[i_ex]:>8s}  {r[i_src].strip()[:90]}")
