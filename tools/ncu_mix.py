"""Instruction mix / hot segments of one kernel from an .ncu-rep source page: python tools/ncu_mix.py rep kernel_regex [dump]"""
import collections
import csv
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{rx}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
# several launches may be concatenated: keep the first
hdr = rows[1]
ia, isrc, ismp = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
body = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name":
        break
    if len(r) > ia:
        body.append(r)
tot = 0
byop, samp = collections.Counter(), collections.Counter()
seq = []
for r in body:
    n = int(r[ia])
    tot += n
    t = r[isrc].split()
    op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    byop[op] += n
    samp[op] += int(r[ismp])
    seq.append((n, r[isrc].strip(), int(r[ismp])))
print("total warp instructions", tot, "samples", sum(s for _, _, s in seq))
for k, v in byop.most_common(28):
    print(f"{k:10s} {v:12d} {100.0 * v / tot:5.1f}%   samples {samp[k]}")
if len(sys.argv) > 3:
    with open(sys.argv[3], "w") as f:
        for i, (n, op, s) in enumerate(seq):
            f.write(f"{i:5d} {n:10d} {s:6d}  {op}\n")
