#!/usr/bin/env python
"""Per-kernel counters that cannot be measured live (DRAM bytes, executed warp instructions, active lanes), per unit of
work, from one `ncu --set full` capture of tools/prof.py plus the bounce trace of the same command:

    MTS_TRACE_BOUNCES=1 python tools/prof.py --spp 8 2> trace.txt          (plain run, same scene / spp)
    ncu --set full --clock-control none -k regex:"k_intersect|k_shade|k_shadow" -c 12 -o cap python tools/prof.py --spp 8
    python tools/ncu_counters.py cap.ncu-rep trace.txt "ncu command line" > profiles/kernel_counters.json

The n-th captured launch of a kernel is matched to bounce n of the first wavefront (units = closest-hit rays for
k_intersect, surface hits for k_shade*, shadow rays for k_shadow).  bench.py multiplies these per-unit figures with its
live unit counts and launch times (`roofline.traffic`, `roofline.issue`); a time taken under ncu is never reported.
Runs where ncu is installed; needs no GPU."""
import csv
import json
import re
import subprocess
import sys


def main():
    rep, trace = sys.argv[1], sys.argv[2]
    cmd = sys.argv[3] if len(sys.argv) > 3 else ""
    bounces = []
    prev_escaped = 0
    for line in open(trace):
        m = re.search(r"wavefront (\d+) bounce (\d+): closest-hit rays (\d+), shadow rays (\d+), continuing (\d+)(?:, escaped so far (\d+))?"
                      r"(?:, shaded lambert/leaf/generic (\d+) (\d+) (\d+))?", line)
        if not m or int(m.group(1)) != 0:
            continue
        rays, srays = int(m.group(3)), int(m.group(4))
        esc = int(m.group(6)) if m.group(6) else prev_escaped
        b = {"intersect": rays, "shadow": srays, "shade": rays - (esc - prev_escaped), "miss": esc - prev_escaped}
        if m.group(7):
            b.update(shade0=int(m.group(7)), shade1=int(m.group(8)), shade2=int(m.group(9)))
        bounces.append(b)
        prev_escaped = esc
    if rep.endswith(".csv"):      # `ncu -i cap.ncu-rep --page raw --csv` already run where the capture was taken
        out = open(rep).read()
    else:
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = [r for r in csv.reader(out.splitlines()) if len(r) > 5]
    h = rows[0]
    seen = {}
    res = {"source": "profiles/kernel_counters.json <- %s" % (cmd or rep), "bounces": bounces}
    per = {}
    for r in rows[2:]:
        d = dict(zip(h, r))
        name = d.get("Kernel Name", "")
        kind = "intersect" if "k_intersect" in name else "shadow" if "k_shadow" in name else "miss" if "k_shade_miss" in name else "shade" if "k_shade" in name else None
        if kind is None:
            continue
        cls = re.search(r"(k_shade\w*)<\(?(?:\w+::)*\w*\)?(\d)>", name) if kind == "shade" else None
        sub = kind + (cls.group(2) if cls else "")      # one k_shade instance per shade class and bounce ...
        stage = cls.group(1) + cls.group(2) if cls else sub   # ... or, staged, k_shade_surface / k_shade_ewa / k_shade_scatter
        n = seen.get(stage, 0)
        seen[stage] = n + 1
        if n >= len(bounces):
            continue
        units = bounces[n].get(sub, bounces[n][kind])
        if units <= 0:
            continue

        def f(k, default=0.0):
            try:
                return float(d.get(k, default) or default)
            except ValueError:
                return default

        def in_bytes(k):
            u = rows[1][h.index(k)] if k in h else "byte"
            scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
            return f(k) * scale

        t_unit = rows[1][h.index("gpu__time_duration.sum")] if "gpu__time_duration.sum" in h else "ms"
        t_ms = f("gpu__time_duration.sum") * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(t_unit, 1.0)
        e = {"kernel": name[:60], "sub": sub, "stage": stage, "bounce": n, "units": units, "ms_under_ncu": t_ms,
             "dram_bytes": in_bytes("dram__bytes_read.sum") + in_bytes("dram__bytes_write.sum"),
             "warp_inst": f("smsp__inst_executed.sum"), "lanes_per_inst": f("smsp__thread_inst_executed_per_inst_executed.ratio"),
             "issue_active_pct": f("smsp__issue_active.avg.pct_of_peak_sustained_active"),
             "warps_active_pct": f("sm__warps_active.avg.pct_of_peak_sustained_active"),
             "l2_hit_pct": f("lts__t_sector_hit_rate.pct"), "l1_hit_pct": f("l1tex__t_sector_hit_rate.pct"),
             "registers": f("launch__registers_per_thread"),
             "stall_long_scoreboard": f("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
             "stall_no_instruction": f("smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio")}
        per.setdefault(kind, []).append(e)
    for kind, key in (("intersect", "k_intersect"), ("shade", "k_shade"), ("shadow", "k_shadow"), ("miss", "k_shade_miss")):
        ls = per.get(kind)
        if not ls:
            continue
        if kind == "shade":   # per-class launches: count every covered bounce's shaded paths once
            cov = sorted({e["bounce"] for e in ls})
            units = sum(bounces[b]["shade"] for b in cov)
            for c in ("shade0", "shade1", "shade2"):
                lc = [e for e in ls if e["sub"] == c]
                if lc:
                    uc = sum(bounces[b].get(c, 0) for b in sorted({e["bounce"] for e in lc}))   # staged kernels share the paths of a bounce
                    res["k_" + c] = {"dram_bytes_per_unit": sum(e["dram_bytes"] for e in lc) / uc, "warp_inst_per_unit": sum(e["warp_inst"] for e in lc) / uc,
                                     "ms_under_ncu": sum(e["ms_under_ncu"] for e in lc), "units_covered": uc, "registers": lc[0]["registers"],
                                     "stall_no_instruction": max(e["stall_no_instruction"] for e in lc), "warps_active_pct": lc[0]["warps_active_pct"]}
        else:
            units = sum(e["units"] for e in ls)
        wi = sum(e["warp_inst"] for e in ls)
        res[key] = {"dram_bytes_per_unit": sum(e["dram_bytes"] for e in ls) / units,
                    "warp_inst_per_unit": wi / units,
                    "lanes_per_inst": sum(e["lanes_per_inst"] * e["warp_inst"] for e in ls) / max(wi, 1),
                    "issue_active_pct": sum(e["issue_active_pct"] * e["ms_under_ncu"] for e in ls) / max(sum(e["ms_under_ncu"] for e in ls), 1e-9),
                    "units_covered": units, "launches": ls}
    json.dump(res, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
