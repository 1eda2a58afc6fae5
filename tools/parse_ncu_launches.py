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
    if "--families" in sys.argv:
        # round-2 form: one step per workload (a step = the launches from one image-stem launch up to the next),
        # DRAM traffic and time share per kernel family -> JSON on stdout (profiles/traffic_rNN.json)
        starts = [i for i, d in enumerate(launches) if d["name"].startswith(("stem_kernel", "stem_tc_kernel"))]
        out = {"source": path, "note": "ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum "
                                        "--clock-control none on `python bench.py --steps 2 --warmup 3`: launches are serialised and "
                                        "cold, read the SHARES; one step = stem launch .. next stem launch"}
        steps_by = {}
        for a, b in zip(starts[:-1], starts[1:]):
            step = launches[a:b]
            wl = "erfnet" if any(x["name"].startswith("conv_pair_tc_kernel") for x in step) else (
                "fastscnn" if any(x["name"].startswith("dw_tma_kernel") for x in step) else None)
            if wl is None or len(step) < 20:
                continue
            steps_by.setdefault(wl, []).append(step)
        seen = {}
        for wl, steps_ in steps_by.items():   # a forward pass has the most common launch count (the roofline pass of
            from collections import Counter   # bench.py re-launches single layers behind the last stem: not a step)
            n_mode = Counter(len(x) for x in steps_).most_common(1)[0][0]
            seen[wl] = [x for x in steps_ if len(x) == n_mode][-1]
        for wl, step in seen.items():
            total = sum(x.get("us", 0.0) for x in step) or 1.0
            fams = {}
            for x in step:
                f = fams.setdefault(x["name"].split("<")[0], {"launches": 0, "us": 0.0, "rd": 0.0, "wr": 0.0})
                f["launches"] += 1
                f["us"] += x.get("us", 0.0)
                f["rd"] += x.get("dram__bytes_read.sum", 0.0)
                f["wr"] += x.get("dram__bytes_write.sum", 0.0)
            out[wl] = {"batch": 16 if wl == "erfnet" else 32, "launches_per_step": len(step), "us_per_step_under_ncu": round(total, 1),
                       "families": {k: {"launches_per_step": v["launches"], "share_of_step": round(v["us"] / total, 4),
                                        "dram_bytes_per_step": int(v["rd"] + v["wr"]), "dram_read_bytes": int(v["rd"]),
                                        "dram_write_bytes": int(v["wr"]),
                                        "note": "dram__bytes_read.sum + dram__bytes_write.sum over this family's launches of one step"}
                                    for k, v in sorted(fams.items(), key=lambda kv: -kv[1]["us"])}}
        print(json.dumps(out, indent=1))
        return
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
