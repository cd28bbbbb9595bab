#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove the hardware path of libfcb200.so (tcgen05 = UTCHMMA, TMEM reads = LDTM,
TMA = UTMALDG, commit / barriers = UTCBAR / SYNCS, shared atomics = ATOMS, global reductions = REDG / ATOMG, 128-bit loads).
  python scripts/sass_summary.py > profiles/r02_sass_summary.txt        (needs cuobjdump; no GPU)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "financial-chatbot-llm_b200", "csrc", "libfcb200.so")
KEYS = ["UTCHMMA", "UTCBAR", "UTMALDG", "LDTM", "UTCATOMSWS", "SYNCS", "ATOMS", "REDG", "ATOMG", "LDG.128", "LDG.64", "LDGSTS", "STG",
        "REDUX", "VOTE", "SHFL", "DFMA", "FFMA", "BAR.SYNC"]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"\n\s*Function : ", txt)
    print("SASS summary of financial-chatbot-llm_b200/csrc/libfcb200.so (cuobjdump -sass, sm_100a): per kernel, how often each")
    print("mnemonic occurs (static count). UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor (TMA),")
    print("UTCBAR = tcgen05.commit, ATOMS = shared-memory atomic, REDG / ATOMG = global reduction / atomic. No UTMALDG carries")
    print(".MULTICAST (TMA multicast loads are not used); UTCBAR.2CTA.MULTICAST is the commit that signals both CTAs of a pair.\n")
    for part in parts[1:]:
        name = part.split("\n", 1)[0].strip()
        c, variants = collections.Counter(), collections.Counter()
        for line in part.split("\n"):
            m = re.search(r"^\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
            if not m:
                continue
            op = m.group(1)
            if op.startswith("LDG"):
                c["LDG.128" if ".128" in op else "LDG.64" if ".64" in op else "LDG"] += 1
                continue
            for k in KEYS:
                if op.startswith(k):
                    c[k] += 1
            if op.startswith(("UTCHMMA", "UTMALDG", "UTCBAR", "LDTM")):
                variants[op] += 1
        if not c:
            continue
        short = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip() or name
        short = re.sub(r"\(.*", "", short).replace("void ", "")
        print(f"{short}: " + ", ".join(f"{k} {c[k]}" for k in KEYS + ["LDG"] if c[k]))
        if variants:
            print("    variants: " + ", ".join(f"{k} x{v}" for k, v in sorted(variants.items())))


if __name__ == "__main__":
    sys.exit(main())
