"""Per-kernel counts of the SASS mnemonics that identify the Blackwell paths (B200_PROFILING.md) in the in-tree library:
python tools/sass_census.py > profiles/sass_census.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "cloth-splatting_b200", "lib", "libcloth_splatting_b200.so")
KEYS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UBLKCP", "UTMALDG", "LDGSTS", "SYNCS", "FFMA2", "FMUL2", "FADD2", "FFMA", "MUFU.EX2",
        "MUFU.RCP", "REDG", "RED.", "ATOMG", "SHFL", "HMMA", "LDL", "STL"]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", txt)))
    counts, name = collections.OrderedDict(), None
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"^void ", "", name).split("(")[0].replace("csb::", "")
            counts[name] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and name:
            op = m.group(1)
            counts[name]["total"] += 1
            for k in KEYS:
                if op == k or op.startswith(k + ".") or (k.endswith(".") and op.startswith(k)) or (("." in k) and op.startswith(k)):
                    counts[name][k] += 1
    print(f"# SASS census of {os.path.relpath(SO, ROOT)} (cuobjdump -sass; arch {', '.join(arch)}): instruction counts per kernel")
    print("# UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk (TMA engine), LDGSTS = cp.async,")
    print("# SYNCS = mbarrier, FFMA2/FMUL2/FADD2 = packed fp32x2, REDG/RED = reduction atomics, LDL/STL = local-memory (spill) traffic")
    used = [k for k in KEYS if any(c[k] for c in counts.values())]
    print("| kernel | total | " + " | ".join(used) + " |")
    print("|---|---|" + "---|" * len(used))
    for n, c in counts.items():
        print(f"| `{n}` | {c['total']} | " + " | ".join(str(c[k]) if c[k] else "." for k in used) + " |")


if __name__ == "__main__":
    sys.exit(main())
