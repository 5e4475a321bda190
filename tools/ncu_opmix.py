"""Executed-instruction mix by opcode of the first kernel in a report: python tools/ncu_opmix.py file.ncu-rep"""
import csv, io, subprocess, sys, collections
txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr=None; data=[]; kern=0
for r in rows:
    if r and r[0]=="Kernel Name": kern+=1; continue
    if r and r[0]=="Address": hdr=r; continue
    if hdr and kern==1 and len(r)>=5: data.append(r)
i_src,i_ex=hdr.index("Source"),hdr.index("Instructions Executed")
mix=collections.Counter()
for r in data:
    toks=r[i_src].split()
    op=toks[1] if toks and toks[0].startswith('@') else (toks[0] if toks else '?')
    mix[op.split('.')[0]]+=int(r[i_ex] or 0)
tot=sum(mix.values())
for op,n in mix.most_common(22): print(f"{op:12s} {n:12d} {100*n/tot:5.1f}%")
print("total", tot)
