#!/usr/bin/env python
"""Instruction mix (SASS opcode histogram) and trimmed listings of selected kernels of libdisb200.so.

  python tools/sass_mix.py of_dis_b200/libdisb200.so profiles/r02_sass

Writes <prefix>_mix.txt (one histogram per kernel; Blackwell-specific mnemonics called out) and <prefix>_<kernel>.sass
(opcode + operands only, no encodings) for the kernels named in HOT.  Needs cuobjdump and c++filt; no GPU.
"""
import collections
import re
import subprocess
import sys

HOT = {"patch_search_8x8_grey_of": "patch_search_kernel<8, 1, 4, true, 0, false, true>", "patch_search_8x8_grey_of_direct": "patch_search_kernel<8, 1, 4, true, 0, false, false>",
       "patch_search_12x12_grey_of_packed": "patch_search_kernel<12, 1, 8, true, 4, true, true>", "sor_warp_of_3sweeps_2bands": "sor_warp_kernel<true, 3, 2>", "sor_team_of_3sweeps_2bands": "sor_team_kernel<true, 3, 2>",
       "system_grey_of": "system_kernel<1, true>", "upsample_of": "upsample_crop_kernel<2>", "sor_band_of_3sweeps": "sor_band_kernel<true, 3>",
       "smooth4_of": "smooth_kernel4<true>", "densify_grey_of": "densify_kernel<1, 2>", "tv_solve_fused_grey_of_2bands": "tv_solve_fused_kernel<1, true, 3, 2>"}
LISTED = ("patch_search_8x8_grey_of", "sor_warp_of_3sweeps_2bands", "sor_team_of_3sweeps_2bands", "system_grey_of")   # full listings
BLACKWELL = ("UTMALDG", "UTMASTG", "UBLKCP", "UTC", "LDTM", "STTM", "SYNCS", "UTMA")
PACKED = ("FFMA2", "FADD2", "FMUL2")


def main():
    lib, prefix = sys.argv[1], sys.argv[2]
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    blocks = re.split(r"\n\s*Function : ", sass)[1:]
    names = subprocess.run(["c++filt"], input="\n".join(b.split("\n", 1)[0].strip() for b in blocks), capture_output=True, text=True).stdout.splitlines()
    total = collections.Counter()
    with open(prefix + "_mix.txt", "w") as mix:
        mix.write("SASS opcode histograms, %s (cuobjdump -sass; tools/sass_mix.py)\n\n" % lib)
        for blk, name in zip(blocks, names):
            ops = collections.Counter()
            lines = []
            for ln in blk.split("\n")[1:]:
                m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", ln)
                if not m:
                    continue
                ins = m.group(1).strip()
                lines.append(ins)
                op = ins.split()[1] if ins.startswith("@") else ins.split()[0]
                ops[op.split(".")[0]] += 1
            total.update(ops)
            for tag, pat in HOT.items():
                if pat in name:
                    n = sum(ops.values())
                    mix.write("%s  [%s]\n  %d instructions: %s\n\n" % (pat, tag, n, ", ".join("%s %d" % kv for kv in ops.most_common(24))))
                    if tag in LISTED:
                        with open("%s_%s.sass" % (prefix, tag), "w") as f:
                            f.write("// %s\n" % name)
                            f.write("\n".join(lines) + "\n")
        bw = {k: v for k, v in total.items() if k.startswith(BLACKWELL)}
        pk = {k: v for k, v in total.items() if k in PACKED}
        mix.write("whole library: %d instructions in %d kernels; async copies: LDGSTS %d (cp.async); packed fp32 pairs (sm_100): %s; bulk / TMA / tcgen05 mnemonics: %s\n"
                  % (sum(total.values()), len(blocks), total.get("LDGSTS", 0), pk if pk else "none",
                     bw if bw else "none (stencil / gather path: no tensor-core or TMA tile traffic)"))


if __name__ == "__main__":
    main()
