#!/usr/bin/env python
"""Per-source-line summary of one kernel from an .ncu-rep captured with --import-source on.
usage: ncu_src.py report.ncu-rep kernel_regex [top_n]"""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
sort_key = sys.argv[4] if len(sys.argv) > 4 else "smp"      # smp | ins
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = []
fname = None; hdr = None
for r in csv.reader(io.StringIO(out)):
    if not r: continue
    if r[0] in ("File Name", "File Path"): fname = r[1].split("/")[-1]; hdr = None; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = ["Line No", "Src"] + r[2:]; continue
    if hdr and r[0] != "" and len(r) >= 10:            # per-source-line aggregate (SASS rows have an empty line number)
        d = dict(zip(hdr, r)); d["file"] = fname; d["Source"] = r[1]; rows.append(d)
def num(d, k):
    try: return float(d.get(k, 0) or 0)
    except ValueError: return 0.0
keyS = "Warp Stall Sampling (All Samples)"
tot = sum(num(d, keyS) for d in rows) or 1
toti = sum(num(d, "Instructions Executed") for d in rows) or 1
print(f"total samples {tot:.0f}, warp instructions {toti:.0f}")
for d in sorted(rows, key=lambda d: -num(d, keyS if sort_key == "smp" else "Instructions Executed"))[:top]:
    print(f"{100*num(d,keyS)/tot:5.1f}% smp {100*num(d,'Instructions Executed')/toti:5.1f}% ins thr/ins {num(d,'Avg. Threads Executed'):4.1f} "
          f"L2sect {num(d,'L2 Theoretical Sectors Global'):10.0f} | {d['file']}:{d['Line No']} {d['Source'].strip()[:110]}")
