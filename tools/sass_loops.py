"""Instruction histogram of every backward-branch loop of a kernel's SASS:
python tools/sass_loops.py <obj.o> <mangled-name-substring>"""
import re, subprocess, sys
from collections import Counter
obj, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
for f in funcs[1:]:
    name = f.split("\n", 1)[0]
    if pat not in name:
        continue
    rows = []
    for l in f.split("\n"):
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)(.*?);", l)
        if m:
            rows.append((int(m.group(1), 16), m.group(2), m.group(3)))
    addr = [r[0] for r in rows]
    print(name, "instructions:", len(rows), "DMMA:", sum(r[1].startswith("DMMA") for r in rows))
    for i, (a, op, rest) in enumerate(rows):
        if op.startswith("BRA"):
            t = re.search(r"0x([0-9a-f]+)", rest)
            if t and int(t.group(1), 16) < a and int(t.group(1), 16) in addr:
                j = addr.index(int(t.group(1), 16))
                c = Counter(r[1].split(".")[0] for r in rows[j:i + 1])
                if c.get("DMMA", 0) or len(rows[j:i+1]) > 40:
                    print(f"  loop {addr[j]:#x}->{a:#x} n={i + 1 - j}", dict(c.most_common(16)))
