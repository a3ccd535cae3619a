#!/usr/bin/env python
"""Top source lines of an ncu report by warp-stall samples:  ncu_lines.py report.ncu-rep [n]
(runs `ncu -i ... --page source --print-source cuda,sass --csv` and aggregates per CUDA source line)."""
import csv, subprocess, sys, io
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
fpath = None; hdr = None; lines = []
for r in csv.reader(io.StringIO(out)):
    if not r: continue
    if r[0] == "File Path": fpath = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or r[0] in ("", "-"): continue
    try:
        si = hdr.index("# Samples"); ie = hdr.index("Instructions Executed")
        rec = {"file": fpath, "line": int(r[0]), "src": r[1].strip()[:100], "samples": int(r[si]), "inst": int(r[ie])}
        for i, h in enumerate(hdr):
            if h.startswith("stall_") and "Not Issued" not in h:
                try: rec[h] = int(r[i])
                except ValueError: rec[h] = 0
        lines.append(rec)
    except (ValueError, IndexError):
        pass
tot = sum(x["samples"] for x in lines) or 1; toti = sum(x["inst"] for x in lines) or 1
print("total samples %d, warp instructions %d" % (tot, toti))
st = {}
for x in lines:
    for k, v in x.items():
        if k.startswith("stall_"): st[k] = st.get(k, 0) + v
print("stalls:", ", ".join("%s %.1f%%" % (k[6:], 100.0 * v / max(1, sum(st.values()))) for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:7]))
for x in sorted(lines, key=lambda x: -x["samples"])[:topn]:
    top = sorted(((v, k[6:]) for k, v in x.items() if k.startswith("stall_")), reverse=True)[:2]
    print("%5.1f%% smp %5.1f%% inst  %-22s:%4d  %-14s %s" % (100.0 * x["samples"] / tot, 100.0 * x["inst"] / toti, x["file"], x["line"],
          "/".join(t[1] for t in top if t[0]), x["src"]))
