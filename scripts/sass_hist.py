"""Opcode histogram of one kernel from `cuobjdump -sass <object>` (evidence for DESIGN.md section 4: which pipes a kernel's
instructions go to).  Usage: python scripts/sass_hist.py <object file> <kernel name substring> [top]"""
import collections
import re
import subprocess
import sys

obj, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True, check=True).stdout
hist, name, total = None, None, 0
res = []
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        if hist is not None:
            res.append((name, hist))
        name = m.group(1)
        hist = collections.Counter() if pat in name else None
        continue
    if hist is None:
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        hist[m.group(1)] += 1
if hist is not None:
    res.append((name, hist))
for name, hist in res:
    total = sum(hist.values())
    demangled = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    print(f"# {demangled}\n# {total} SASS instructions")
    groups = collections.Counter()
    for op, n in hist.items():
        base = op.split(".")[0]
        groups["IMAD.WIDE*" if op.startswith("IMAD.WIDE") else base] += n
    print("# by class:", ", ".join(f"{k} {v}" for k, v in groups.most_common(14)))
    for op, n in hist.most_common(top):
        print(f"{n:7d}  {100.0 * n / total:5.1f}%  {op}")
    print()
