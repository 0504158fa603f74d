#!/usr/bin/env python
"""Instruction / stall-sample shares of one kernel by source file and by line ranges of one file, from an .ncu-rep captured with
--import-source on.   usage: ncu_agg.py report.ncu-rep kernel_regex [file [lines_per_bucket]]"""
import csv, io, subprocess, sys, collections
rep, kern = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
rows = []; fname = None; hdr = None
for r in csv.reader(io.StringIO(out)):
    if not r: continue
    if r[0] in ("File Name", "File Path"): fname = r[1].split("/")[-1]; hdr = None; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = ["Line No", "Src"] + r[2:]; continue
    if hdr and r[0] != "" and len(r) >= 10:
        d = dict(zip(hdr, r)); d["file"] = fname; rows.append(d)
def num(d, k):
    try: return float(d.get(k, 0) or 0)
    except ValueError: return 0.0
byf = collections.Counter(); bys = collections.Counter()
for d in rows:
    byf[d["file"]] += num(d, "Instructions Executed"); bys[d["file"]] += num(d, "Warp Stall Sampling (All Samples)")
ti = sum(byf.values()) or 1; ts = sum(bys.values()) or 1
for f, v in byf.most_common(): print(f"{f:28s} ins {100*v/ti:5.1f}%  smp {100*bys[f]/ts:5.1f}%")
if len(sys.argv) > 3:
    f = sys.argv[3]; step = int(sys.argv[4]) if len(sys.argv) > 4 else 20
    b = collections.Counter(); s = collections.Counter()
    for d in rows:
        if d["file"] == f:
            k = int(d["Line No"]) // step * step; b[k] += num(d, "Instructions Executed"); s[k] += num(d, "Warp Stall Sampling (All Samples)")
    for k in sorted(b): print(f"  {f}:{k:4d}-{k+step-1:4d} ins {100*b[k]/ti:5.1f}% smp {100*s[k]/ts:5.1f}%")
