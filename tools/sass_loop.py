"""Developer helper: size and opcode mix of the phase-1 loop of the P3M pair kernel in a compiled object
(cuobjdump -sass, no GPU needed): the basic block that holds the eight predicated STS.U16 pushes.
    python tools/sass_loop.py nbody_b200/csrc/p3m.o [kernel-name-substring]"""
import collections
import re
import subprocess
import sys

obj = sys.argv[1]
want = sys.argv[2] if len(sys.argv) > 2 else "p3m_short_v6"
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
for fn in txt.split("Function : ")[1:]:
    name = fn.split("\n", 1)[0]
    if want not in name:
        continue
    ins = [m.group(1).strip() for m in re.finditer(r"/\*[0-9a-f]{4}\*/\s+(.*?);", fn)]
    # split into blocks at branches
    blocks, cur = [], []
    for i in ins:
        cur.append(i)
        if re.search(r"\bBRA\b|\bEXIT\b|\bRET\b", i):
            blocks.append(cur)
            cur = []
    best = None
    for k, b in enumerate(blocks):
        n = sum("STS.U16" in x for x in b)
        if n >= 8 and sum("LDS.128" in x for x in b) + sum("LDS.128" in x for x in (blocks[k - 1] if k else [])) >= 6:
            best = k
            break
    print(name[:90])
    if best is None:
        print("  phase-1 block not found")
        continue
    # the loop = this block plus the short continuation blocks up to the back edge
    loop = list(blocks[best])
    j = best + 1
    while j < len(blocks) and len(blocks[j]) < 12 and "STS.U16" not in " ".join(blocks[j]):
        loop += blocks[j]
        j += 1
    ops = collections.Counter((x.split()[1] if x.startswith("@") else x.split()[0]).split(".")[0] for x in loop)
    print(f"  loop instructions: {len(loop)}   {dict(ops.most_common())}")
