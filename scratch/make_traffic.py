"""profiles/r01_traffic.json from the two ncu --set full raw pages: DRAM bytes per pair, per kernel group."""
import csv, json, collections, sys
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
GROUP = [("sor_", "sor"), ("downscale", "pyramid"), ("halve", "pyramid"), ("sobel", "pyramid"), ("patch_search", "inverse_search"),
         ("densify", "densify"), ("bin_", "densify"), ("warp_kernel", "varref_prep"), ("deriv", "varref_prep"), ("finish", "varref_prep"),
         ("smooth", "smooth"), ("system", "system"), ("upsample", "upsample")]
def load(path, batch, keep):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    ir, iw, ik = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
    acc = collections.defaultdict(lambda: [0.0, 0.0, 0])
    seen_ps = 0
    for r in rows[2:]:
        if len(r) <= iw: continue
        name = r[ik]
        g = next((grp for key, grp in GROUP if key in name), None)
        if g is None: continue
        if not keep(name, r, hdr): continue
        acc[g][0] += float(r[ir].replace(",", "")) * UNIT[units[ir]] / batch
        acc[g][1] += float(r[iw].replace(",", "")) * UNIT[units[iw]] / batch
        acc[g][2] += 1
    return acc
root = sys.argv[1] if len(sys.argv) > 1 else "."
# r01c: batch 1480, pyramid + levels 5 and 4 (the capture ends inside level 3: drop its lone level-3 patch search)
def keep_c(name, r, hdr):
    return not ("patch_search" in name and r[hdr.index("launch__grid_size")].replace(",", "") == "41440")
a = load(root + "/profiles/r01c_step_full_raw.csv", 1480, keep_c)
b = load(root + "/profiles/r01f_l3_full_raw.csv", 592, lambda *x: True)
out = {}
for g in sorted(set(a) | set(b)):
    out[g] = {"dram_read_bytes_per_pair": round(a[g][0] + b[g][0]), "dram_write_bytes_per_pair": round(a[g][1] + b[g][1]),
              "launches": a[g][2] + b[g][2]}
json.dump({"source": "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch, summed per kernel group and divided by the pairs of the capture: level 3 + upsample from r01f_l3_full_raw.csv (batch 592), pyramid + levels 5, 4 from r01c_step_full_raw.csv (batch 1480); built by scratch/make_traffic.py",
           "workload": "C1", "per_stage": out}, open(root + "/profiles/r01_traffic.json", "w"), indent=1)
print(json.dumps(out, indent=1))
