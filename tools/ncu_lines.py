"""Per-source-line summary of an ncu report (SourceCounters): python tools/ncu_lines.py rep.ncu-rep [bins] [top]"""
import csv, subprocess, sys
from collections import defaultdict
rep = sys.argv[1]; bins = float(sys.argv[2]) if len(sys.argv) > 2 else 0; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
cur = None; hdr = None
agg = defaultdict(lambda: [0.0, 0.0, ""]); stall = defaultdict(float)
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": cur = r[1].split('/')[-1]; continue
    if len(r) >= 2 and r[0] == "Function Name": continue
    if r and r[0] == "Line No":
        hdr = r; iS = hdr.index("# Samples"); iI = hdr.index("Instructions Executed")
        st = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        continue
    if hdr is None or len(r) < len(hdr): continue
    try: ln = int(r[0])
    except ValueError: continue
    try: s = float(r[iS] or 0); i = float(r[iI] or 0)
    except ValueError: continue
    k = (cur, ln); agg[k][0] += s; agg[k][1] += i
    if r[1]: agg[k][2] = r[1]
    for j, h in st:
        try: stall[h] += float(r[j] or 0)
        except ValueError: pass
ts = sum(v[0] for v in agg.values()); ti = sum(v[1] for v in agg.values())
print(f"samples {ts:.0f}  instructions {ti:.0f}" + (f"  instr/bin {ti / bins:.1f}" if bins else ""))
for h, v in sorted(stall.items(), key=lambda x: -x[1])[:8]: print(f"  {h:26s}{100 * v / max(ts, 1):5.1f}%")
byf = defaultdict(lambda: [0.0, 0.0])
for (f, l), v in agg.items(): byf[f][0] += v[0]; byf[f][1] += v[1]
for f, v in sorted(byf.items(), key=lambda x: -x[1][0]): print(f"  {f:24s}{100 * v[0] / ts:5.1f}% samples {100 * v[1] / ti:5.1f}% instr" + (f" {v[1] / bins:6.1f} i/bin" if bins else ""))
for (f, l), v in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print(f"{f}:{l:4d} {100 * v[0] / ts:5.1f}% s {100 * v[1] / ti:5.1f}% i" + (f" {v[1] / bins:5.1f}/bin" if bins else "") + f"  {v[2][:100]}")
