#!/usr/bin/env python
"""Opcode inventory of the built library: which SASS tensor / TMA / TMEM instructions each kernel contains.

  python tools/sass_inventory.py > profiles/r02_sass_inventory.txt

UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG = TMA tile load / store, UBLKCP = bulk copy,
HMMA = legacy mma.sync (only the fp32-accurate encoder attention, <2 % of that chain's FLOPs, uses it), LDGSTS = cp.async.
"""
import re
import subprocess
import sys
from collections import Counter, OrderedDict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "transformer_tts_b200" / "csrc" / "libttsb200.so"
OPS = ["UTCHMMA", "UTCQMMA", "UTCIMMA", "UTCOMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "UTCBAR",
       "UTCCP", "SYNCS", "HMMA", "HGMMA", "LDGSTS", "MUFU", "FFMA2", "FADD2", "REDG", "ATOMG", "STG", "LDG", "STS", "LDS"]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True).stdout
    kernels: "OrderedDict[str, Counter]" = OrderedDict()
    cur = None
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = kernels.setdefault(re.sub(r"\(.*", "", name), Counter())
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)(\.[A-Z0-9_.]+)?", line)
        if m:
            op, suffix = m.group(1), m.group(2) or ""
            cur[op] += 1
            cur["_total"] += 1
            if op.startswith("UTC") and "2CTA" in suffix:
                cur[op + ".2CTA"] += 1
    git = subprocess.run(["git", "-C", str(ROOT), "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    print(f"# cuobjdump -sass {LIB.relative_to(ROOT)} (sm_100a), tree {git}; instruction counts per kernel")
    tot = Counter()
    for k, c in kernels.items():
        tot.update(c)
        keys = [o for o in OPS if c.get(o)] + [o for o in c if o.endswith(".2CTA")]
        print(f"{k}\n    total {c['_total']:6d}  " + "  ".join(f"{o} {c[o]}" for o in keys))
    print("\n# library totals: " + "  ".join(f"{o} {tot[o]}" for o in OPS + [o for o in tot if o.endswith('.2CTA')] if tot.get(o)))


if __name__ == "__main__":
    main()
