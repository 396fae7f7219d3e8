#!/usr/bin/env python
"""Per-source-line share of executed warp instructions and stall samples from an ncu report
(`ncu --set full --import-source on`): python tools/ncu_lines.py report.ncu-rep [min_pct]."""
import csv, subprocess, sys
rep = sys.argv[1]; thr = float(sys.argv[2]) if len(sys.argv) > 2 else 1.5
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
kern = hdr = f = None; data = {}
for r in csv.reader(txt.splitlines()):
    if len(r) == 2 and r[0] == "File Path": f = r[1].split("/")[-1]; continue
    if len(r) == 2 and r[0] == "Function Name": kern = r[1][:70]; data.setdefault(kern, []); continue
    if r and r[0] == "Line No": hdr = r; continue
    if kern and hdr and len(r) == len(hdr) and r[0] != "": data[kern].append((f, r))
ix = {n: i for i, n in enumerate(hdr)}
for k, v in data.items():
    tot = sum(int(r[ix["Instructions Executed"]]) for _, r in v); samp = sum(int(r[ix["# Samples"]]) for _, r in v)
    print("==", k, "warp inst", tot, "samples", samp)
    for f, r in v:
        ie = int(r[ix["Instructions Executed"]]); s = int(r[ix["# Samples"]])
        if ie > thr / 100 * tot or s > thr / 100 * samp:
            print(f'{f[:14]:14} {r[0]:>4} inst {100*ie/tot:5.1f}% samp {100*s/samp:5.1f}% shExc {r[ix["L1 Wavefronts Shared Excessive"]]:>9} | {r[1].strip()[:100]}')
