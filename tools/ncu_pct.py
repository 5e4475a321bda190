"""All *pct_of_peak* style metrics above a threshold for the first launch in a report: python tools/ncu_pct.py file.ncu-rep [min_pct]"""
import csv, io, subprocess, sys
txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units, r = rows[0], rows[1], rows[2]
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 30.0
for h, u, v in zip(hdr, units, r):
    if u == "%" :
        try: x = float(v.replace(",", ""))
        except ValueError: continue
        if x >= thr: print(f"{x:8.2f}  {h}")
