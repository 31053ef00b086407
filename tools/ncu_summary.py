#!/usr/bin/env python
"""Digest of an `ncu --page raw --csv` export: one record per kernel group (stage of the path).

  ncu -i gpurun_out/r02h_step.ncu-rep --page raw --csv > profiles/r02h_step_full_raw.csv
  python tools/ncu_summary.py profiles/r02h_step_full_raw.csv <pairs in the capture> profiles/r02_kernels.json

Per group: launches, summed duration, DRAM bytes read / written per pair (dram__bytes_*.sum), the duration-weighted means of
issue-slot utilisation (smsp__issue_active), achieved occupancy (sm__warps_active), DRAM throughput (% of peak), L1 hit rate,
registers per thread, executed warp instructions.  bench.py reads `per_stage[*].dram_*_bytes_per_pair` (roofline.traffic) and
`issue_active_pct` (the issue-slot roofline of the instruction-bound stages) from the JSON.
"""
import collections
import csv
import json
import sys

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3, "second": 1.0,
        "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}
GROUP = [("sor_", "sor"), ("tv_solve", "sor"), ("tv_fused", "sor"), ("tv_narrow", "sor"), ("downscale", "pyramid"), ("halve", "pyramid"), ("sobel", "pyramid"),
         ("patch_search", "inverse_search"), ("densify", "densify"), ("bin_", "densify"), ("warp_kernel", "varref_prep"),
         ("deriv", "varref_prep"), ("finish", "varref_prep"), ("smooth", "smooth"), ("system", "system"), ("upsample", "upsample"),
         ("crop_kernel", "upsample"), ("warm_init", "upsample")]
COLS = {"dur": "gpu__time_duration.sum", "rd": "dram__bytes_read.sum", "wr": "dram__bytes_write.sum",
        "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active", "occ": "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1hit": "l1tex__t_sector_hit_rate.pct",
        "regs": "launch__registers_per_thread", "inst": "smsp__inst_executed.sum", "l2hit": "lts__t_sector_hit_rate.pct"}


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return None


def main():
    path, pairs, out = sys.argv[1], float(sys.argv[2]), sys.argv[3]
    rows = list(csv.reader(open(path)))
    h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units = rows[h], rows[h + 1]
    ik = hdr.index("Kernel Name")
    idx = {k: hdr.index(v) for k, v in COLS.items() if v in hdr}
    acc = collections.OrderedDict()
    per_kernel = collections.OrderedDict()
    for r in rows[h + 2:]:
        if len(r) <= ik:
            continue
        name = r[ik]
        g = next((grp for key, grp in GROUP if key in name), None)
        if g is None:
            continue
        v = {k: num(r[i]) for k, i in idx.items()}
        dur = (v.get("dur") or 0.0) * UNIT.get(units[idx["dur"]], 1e-9) if "dur" in idx else 0.0
        rd = (v.get("rd") or 0.0) * UNIT.get(units[idx["rd"]], 1) if "rd" in idx else 0.0
        wr = (v.get("wr") or 0.0) * UNIT.get(units[idx["wr"]], 1) if "wr" in idx else 0.0
        short = name.split("(")[0].replace("void ", "").replace("disb200::", "").replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
        for key, table in ((g, acc), (short, per_kernel)):
            a = table.setdefault(key, {"launches": 0, "seconds": 0.0, "rd": 0.0, "wr": 0.0, "w": collections.defaultdict(float), "regs": 0, "inst": 0.0})
            a["launches"] += 1
            a["seconds"] += dur
            a["rd"] += rd
            a["wr"] += wr
            for k in ("issue", "occ", "dram_pct", "l1hit", "l2hit"):
                if v.get(k) is not None:
                    a["w"][k] += v[k] * dur
            a["regs"] = max(a["regs"], int(v.get("regs") or 0))
            a["inst"] += v.get("inst") or 0.0
    total = sum(a["seconds"] for a in acc.values())

    def fin(a):
        d = {"launches": a["launches"], "ms": round(a["seconds"] * 1e3, 4), "share_of_serialised_step": round(a["seconds"] / total, 4) if total else None,
             "dram_read_bytes_per_pair": round(a["rd"] / pairs), "dram_write_bytes_per_pair": round(a["wr"] / pairs),
             "dram_gbs": round((a["rd"] + a["wr"]) / a["seconds"] / 1e9, 1) if a["seconds"] else None,
             "registers_per_thread_max": a["regs"], "warp_instructions_per_pair": round(a["inst"] / pairs)}
        for k, nm in (("issue", "issue_active_pct"), ("occ", "achieved_occupancy_pct"), ("dram_pct", "dram_throughput_pct"), ("l1hit", "l1_hit_pct"), ("l2hit", "l2_hit_pct")):
            if a["seconds"] and k in a["w"]:
                d[nm] = round(a["w"][k] / a["seconds"], 2)
        return d
    res = {"source": "%s (ncu --set full --clock-control none, per-launch values; sums per group, percentages weighted by launch duration), %g pairs in the capture" % (path, pairs),
           "per_stage": {k: fin(a) for k, a in acc.items()}, "per_kernel": {k: fin(a) for k, a in per_kernel.items()}}
    with open(out, "w") as f:
        json.dump(res, f, indent=1)
    for k, d in res["per_stage"].items():
        print(k, d)


if __name__ == "__main__":
    main()
