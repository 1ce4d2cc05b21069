#!/usr/bin/env python
"""SASS check used while tuning the density kernels: for every loop (backward branch) of a kernel that contains DMMA,
count the DMMA / DFMA / LDS / local-memory (spill) instructions inside it.
usage: sass_loops.py file.o kernel_name_substring"""
import re, subprocess, sys
obj, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout.splitlines()
funcs, cur = {}, None
for l in txt:
    if "Function :" in l:
        cur = l.split(":", 1)[1].strip(); funcs[cur] = []
    elif cur is not None:
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            funcs[cur].append((int(m.group(1), 16), m.group(2).strip()))
for name, ins in funcs.items():
    if pat not in name:
        continue
    print(name[:100], len(ins), "instructions")
    loops = []
    for addr, op in ins:
        m = re.search(r"BRA(?:\.\w+)*\s+(?:\w+,\s*)?`?\(?\.?L?_?x?_?\w*\)?", op)
        if op.startswith("BRA") or " BRA" in op:
            t = re.search(r"0x([0-9a-f]+)", op)
            if t and int(t.group(1), 16) < addr:
                loops.append((int(t.group(1), 16), addr))
    loops = sorted(set(loops))
    inner = [(lo, hi) for lo, hi in loops if not any((l2, h2) != (lo, hi) and lo <= l2 and h2 <= hi and
             any("DMMA" in op for a, op in ins if l2 <= a <= h2) for l2, h2 in loops)]
    for lo, hi in inner:
        body = [op for a, op in ins if lo <= a <= hi]
        n = lambda k: sum(1 for op in body if k in op)
        if n("DMMA"):
            print(f"  loop {lo:#x}-{hi:#x}: {len(body)} instr, DMMA {n('DMMA')}, DFMA {n('DFMA')}, LDS {n('LDS')}, STL {n('STL')}, LDL {n('LDL')}, SYNCS {n('SYNCS')}")
