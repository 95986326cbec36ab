"""Compact reading of an `ncu --page source --csv` export: per phase (split at BAR.SYNC) instructions, samples, shared
wavefronts and the top stall sites.  Usage: python tools/ncu_src_summary.py src.csv n_nodes"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
N = float(sys.argv[2])
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
phase, ph, top = 0, collections.defaultdict(lambda: collections.Counter()), []
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    try:
        e = int(r[ix["Instructions Executed"]] or 0)
    except ValueError:
        break                                                 # second kernel of the export
    s = r[ix["Source"]]
    smp = int(r[ix["# Samples"]] or 0)
    c = ph[phase]
    c["inst"] += e
    c["samples"] += smp
    c["smem_wf"] += int(r[ix["L1 Wavefronts Shared"]] or 0)
    c["gl_req"] += int(r[ix["L1 Tag Requests Global"]] or 0)
    for st in stalls:
        c[st] += int(r[ix[st]] or 0)
    top.append((smp, phase, s[:90], {st: int(r[ix[st]] or 0) for st in stalls if int(r[ix[st]] or 0) > smp * 0.3}))
    if "BAR.SYNC" in s:
        phase += 1
for p in sorted(ph):
    c = ph[p]
    print(f"phase {p}: inst/node {c['inst'] / N:.1f} samples {c['samples']} smem wavefronts/node {c['smem_wf'] / N:.1f} "
          f"global requests/node {c['gl_req'] / N:.1f}")
    print("   stalls:", ", ".join(f"{k[6:]} {v}" for k, v in sorted(((k, c[k]) for k in stalls), key=lambda kv: -kv[1])[:7]))
print("top stall sites:")
for smp, p, s, why in sorted(top, key=lambda t: -t[0])[:28]:
    print(f"  {smp:5d} ph{p} {s}  {why}")
