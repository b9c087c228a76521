"""Instruction histogram of the built library (cuobjdump -sass), whole library and per kernel family, to show which
Blackwell instructions the kernels use (UTCHMMA = tcgen05.mma, UTMALDG / UTMASTG = TMA, LDTM = tcgen05.ld, ...).

  python tools/sass_histogram.py > profiles/r02_sass_histogram.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "medical-image-analysis_b200", "b200al", "libb200al.so")
INS = re.compile(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)")
FAMILIES = ["conv_tc_kernel", "conv_fold_kernel", "conv_first_kernel", "uncertainty_logits_kernel", "dice_logits_kernel",
            "head_uncertainty_kernel", "kcenter_persist_kernel", "kcenter_step_kernel", "kcenter_init_kernel", "topk_",
            "kmpp_", "badge_", "prep_"]
SHOW = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMASTG", "UTMAPF", "LDTM", "SYNCS", "UTCATOM", "HMMA", "DFMA", "DADD", "F2F",
        "MUFU", "SHFL", "LDG", "STG", "LDS", "STS", "FFMA", "ATOM", "RED", "BAR", "ELECT"]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    total, per, kernels = collections.Counter(), collections.defaultdict(collections.Counter), collections.Counter()
    fam = None
    for line in txt.splitlines():
        if "Function :" in line:
            name = line.split("Function :")[1].strip()
            fam = next((f for f in FAMILIES if f in name), "other")
            kernels[fam] += 1
            continue
        m = INS.match(line)
        if m:
            total[m.group(1)] += 1
            per[fam][m.group(1)] += 1
    print(f"# {os.path.relpath(LIB, ROOT)}: {sum(kernels.values())} kernels, {sum(total.values())} SASS instructions")
    print("# whole library, instructions with a modifier-qualified opcode, most frequent first (top 60)")
    for k, v in total.most_common(60):
        print(f"{v:8d}  {k}")
    print("\n# per kernel family: kernels, instructions, and the Blackwell / memory / math opcodes (base mnemonic)")
    for f in FAMILIES + ["other"]:
        if not kernels[f]:
            continue
        base = collections.Counter()
        for k, v in per[f].items():
            base[k.split(".")[0]] += v
        picks = "  ".join(f"{b}={base[b]}" for b in SHOW if base[b])
        print(f"{f:28s} kernels={kernels[f]:3d} instr={sum(per[f].values()):7d}  {picks}")
    two_cta = sum(v for k, v in total.items() if k.startswith("UTCHMMA") and "2CTA" in k)
    print(f"\n# UTCHMMA total {sum(v for k, v in total.items() if k.startswith('UTCHMMA'))} (of which .2CTA {two_cta}); "
          f"UTMALDG {sum(v for k, v in total.items() if k.startswith('UTMALDG'))}; UTMASTG "
          f"{sum(v for k, v in total.items() if k.startswith('UTMASTG'))}; LDTM {sum(v for k, v in total.items() if k.startswith('LDTM'))}")


if __name__ == "__main__":
    main()
