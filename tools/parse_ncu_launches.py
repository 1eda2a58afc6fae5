#!/usr/bin/env python
"""Summarise an `ncu --csv` launch list (gpu__time_duration.sum [+ dram__bytes_*]) per kernel.

    python tools/parse_ncu_launches.py gpurun_out/launches.csv [--steps K] > profiles/launches_rNN.md

Also writes <csv>.json with the per-step DRAM traffic of the conv_tc family (what bench.py reports as
roofline.traffic).  ncu serialises launches and runs them cold: compare SHARES, not absolute times.
"""
from __future__ import annotations

import csv
import json
import re
import sys
from collections import OrderedDict


def short(name: str) -> str:
    name = re.sub(r"^void\s+", "", name)
    name = re.sub(r"\(.*$", "", name)
    for junk in ("rtseg::", "<unnamed>::", "unnamed>::", "(anonymous namespace)::"):
        name = name.replace(junk, "")
    return name


def main():
    path = sys.argv[1]
    steps = int(sys.argv[sys.argv.index("--steps") + 1]) if "--steps" in sys.argv else 1
    rows = [r for r in csv.reader(open(path, errors="replace")) if r]
    hdr_i = next(i for i, r in enumerate(rows) if "Kernel Name" in r and "Metric Name" in r)
    hdr = rows[hdr_i]
    col = {h: i for i, h in enumerate(hdr)}
    per_launch = OrderedDict()
    for r in rows[hdr_i + 1:]:
        if len(r) < len(hdr):
            continue
        key = r[col["ID"]]
        d = per_launch.setdefault(key, {"name": short(r[col["Kernel Name"]])})
        try:
            val = float(r[col["Metric Value"]].replace(",", ""))
        except ValueError:
            continue
        unit = r[col["Metric Unit"]]
        metric = r[col["Metric Name"]]
        if metric == "gpu__time_duration.sum":
            scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit.replace("second", "s").replace("usecond", "us"), None)
            if scale is None:
                scale = {"nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6}.get(unit, 1e-3)
            d["us"] = val * scale
        elif metric.startswith("dram__bytes"):
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
            d[metric] = val * scale
    launches = list(per_launch.values())
    if "--one-step" in sys.argv:
        # one forward pass = from a stem kernel launch up to (not including) the next one
        starts = [i for i, d in enumerate(launches) if d["name"].startswith(("stem_kernel", "stem_tc_kernel"))]
        k = min(int(sys.argv[sys.argv.index("--one-step") + 1]), len(starts) - 2)
        launches = launches[starts[k]:starts[k + 1]]
        steps = 1
    agg = OrderedDict()
    for d in launches:
        a = agg.setdefault(d["name"], {"n": 0, "us": 0.0, "rd": 0.0, "wr": 0.0})
        a["n"] += 1
        a["us"] += d.get("us", 0.0)
        a["rd"] += d.get("dram__bytes_read.sum", 0.0)
        a["wr"] += d.get("dram__bytes_write.sum", 0.0)
    total = sum(a["us"] for a in agg.values()) or 1.0
    print(f"| kernel | launches | total us | share | DRAM read MB | DRAM write MB |")
    print("|---|---:|---:|---:|---:|---:|")
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1]["us"]):
        print(f"| `{name}` | {a['n']} | {a['us']:.1f} | {100 * a['us'] / total:.1f} % | {a['rd'] / 1e6:.1f} | {a['wr'] / 1e6:.1f} |")
    print(f"\n{len(launches)} launches, {total:.1f} us in total ({steps} captured step(s)).")
    fam = [a for n, a in agg.items() if n.startswith("conv_tc_kernel")]
    out = {"steps": steps,
           "conv_tc_launches_per_step": sum(a["n"] for a in fam) / steps,
           "conv_tc_dram_bytes_per_step": sum(a["rd"] + a["wr"] for a in fam) / steps,
           "conv_tc_share_of_step": sum(a["us"] for a in fam) / total,
           "all_dram_bytes_per_step": sum(a["rd"] + a["wr"] for a in agg.values()) / steps}
    json.dump(out, open(path + ".json", "w"), indent=1)


if __name__ == "__main__":
    main()
