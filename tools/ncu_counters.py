"""Collect the ncu counter CSVs of tools/gpu_r2_ncu.sh into profiles/r2_ncu_counters.json (what bench.py's
`roofline.traffic` reads) and print a table.   python tools/ncu_counters.py gpurun_out"""
import csv, json, os, sys
src = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
out = {}
for cfg in (3, 4, 5):
    path = os.path.join(src, f"counters_c{cfg}.csv")
    if not os.path.exists(path):
        continue
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    ki, mi, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    idi = hdr.index("ID")
    per = {}
    for r in rows[1:]:
        try:
            val = float(r[vi].replace(",", ""))
        except ValueError:
            continue  # "n/a": the counter does not exist on this chip
        per.setdefault(r[idi], {"kernel": r[ki]})[r[mi]] = val
        per[r[idi]].setdefault("_units", {})[r[mi]] = r[ui]
    # config 5 runs two launches (one per connected component): sum the additive counters
    ks = list(per.values())
    e = {"kernels": [k["kernel"] for k in ks], "launches_summed": len(ks)}
    for m in ks[0]:
        if m in ("kernel", "_units"):
            continue
        vals = [k[m] for k in ks if m in k]
        additive = m.endswith(".sum")
        e[m] = sum(vals) if additive else max(vals)
    e["units"] = ks[0]["_units"]
    out[f"config{cfg}"] = e
json.dump(out, open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "r2_ncu_counters.json"), "w"), indent=1)
for k, e in out.items():
    print(k, e["kernels"][0][:90])
    for m, v in e.items():
        if m not in ("kernels", "units", "launches_summed"):
            print(f"   {m:70s} {v:,.2f} {e['units'].get(m, '')}")
