"""SASS opcode counts per kernel of the built library (profiles/sass_r02.md):  python tools/sass_counts.py > profiles/sass_r02.md"""
import re
import subprocess
import sys
from collections import Counter, OrderedDict
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
LIB = ROOT / "masters-thesis_b200" / "libb200lstm.so"
COLS = ["UTCHMMA", "LDTM", "UTCBAR", "UBLKCP", "UBLKPF", "UTMALDG", "UTMASTG", "STTM", "SYNCS", "MUFU", "FENCE"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
    kernels, cur, i = OrderedDict(), None, 0
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            full = names[i]; i += 1
            short = full.replace("(anonymous namespace)::", "").replace("b200::", "").replace("tc::", "")
            short = re.sub(r"^void ", "", re.sub(r"\(.*", "", short))
            cur = kernels.setdefault(short, Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur is not None:
            cur[m.group(1)] += 1
            cur["total"] += 1
    print("# SASS opcode counts of the shipped libb200lstm.so (round 2 final; `python tools/sass_counts.py`, per kernel; kernels without a")
    print("# tensor-core, bulk-copy or mbarrier instruction are left out)\n")
    print("| kernel | " + " | ".join(COLS) + " | total |")
    print("|---|" + "---|" * (len(COLS) + 1))
    for k, c in kernels.items():
        if not any(c[x] for x in COLS[:8] + ["SYNCS"]):
            continue
        print(f"| `{k}` | " + " | ".join(str(c[x]) for x in COLS) + f" | {c['total']} |")
    mc = len(re.findall(r"UBLKCP\S*MULTICAST", sass))
    print(f"\nMulticast bulk copies (`UBLKCP...MULTICAST`): {mc} instructions, all in `w2_fwd_kernel<true>` (option `wide_mc`, off by default; "
          "sources are activation images of this GPU's own workspace).  `UTMALDG` / `UTMASTG` (tensor-map TMA) are not used: every operand image "
          "is contiguous, so the 1-D bulk copy (`UBLKCP`) moves it.")


if __name__ == "__main__":
    sys.exit(main())
