"""Summarise ncu artefacts brought back in gpurun_out/ (runs on the CPU box: `ncu -i` needs no GPU).
usage: python tools/ncu_summary.py <tag> [out.txt]      reads gpurun_out/<tag>_launches.csv and gpurun_out/<tag>_*.ncu-rep"""
import csv, glob, io, os, subprocess, sys
from collections import OrderedDict

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_fp64.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"]


def to_bytes(v, unit):
    m = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return float(v.replace(",", "")) * m.get(unit, 1)


def launches(path, out):
    rows = [r for r in csv.reader(l for l in open(path) if not l.startswith("=="))]
    hdr = rows[0]; ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = OrderedDict()
    for r in rows[1:]:
        if len(r) <= iv: continue
        name = r[ik].split("(")[0].split("<")[0]
        t = float(r[iv].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[iu], 1e-3)
        a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += t
    tot = sum(a[1] for a in agg.values())
    print(f"# launch list {os.path.basename(path)}: {sum(a[0] for a in agg.values())} launches, {tot/1e3:.3f} ms under ncu (cold-cache, serialised: compare SHARES)", file=out)
    print(f"{'kernel':40s} {'n':>6s} {'total us':>12s} {'avg us':>10s} {'share':>7s}", file=out)
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:40s} {n:6d} {t:12.1f} {t/n:10.1f} {100*t/tot:6.1f}%", file=out)


def report(path, out):
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    if len(rows) < 3: print(f"# {path}: unreadable", file=out); return
    hdr, units = rows[0], rows[1]
    print(f"\n# {os.path.basename(path)}", file=out)
    for r in rows[2:]:
        d = dict(zip(hdr, r)); u = dict(zip(hdr, units))
        print(f"## {d.get('Kernel Name','?')[:90]}  grid {d.get('Grid Size')} block {d.get('Block Size')}", file=out)
        rd = to_bytes(d["dram__bytes_read.sum"], u["dram__bytes_read.sum"]); wr = to_bytes(d["dram__bytes_write.sum"], u["dram__bytes_write.sum"])
        tns = float(d["gpu__time_duration.sum"].replace(",", "")) * {"ns": 1, "us": 1e3, "ms": 1e6}.get(u["gpu__time_duration.sum"], 1)
        print(f"   dram read {rd/1e6:.2f} MB + write {wr/1e6:.2f} MB = {(rd+wr)/1e6:.2f} MB per launch; duration {tns/1e3:.1f} us (under ncu) -> {(rd+wr)/tns:.1f} GB/s", file=out)
        for k in KEYS:
            if k in d and k not in ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"):
                print(f"   {k} = {d[k]} {u.get(k,'')}", file=out)


def main():
    tag = sys.argv[1]
    out = open(sys.argv[2], "w") if len(sys.argv) > 2 else sys.stdout
    lp = f"gpurun_out/{tag}_launches.csv"
    if os.path.exists(lp): launches(lp, out)
    for p in sorted(glob.glob(f"gpurun_out/{tag}_*.ncu-rep")): report(p, out)


if __name__ == "__main__":
    main()
