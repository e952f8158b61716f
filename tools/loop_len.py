#!/usr/bin/env python
"""Length (SASS instructions) of the innermost loops of a kernel in volume-rendering_b200/build/vr5_render.o: every backward branch
whose body contains a given mnemonic (default TLD or LDG.E.64: the march loops). Usage: loop_len.py <mangled-substring> [mnemonic]"""
import re, subprocess, sys
obj = sys.argv[3] if len(sys.argv) > 3 else "volume-rendering_b200/build/vr5_render.o"
names = subprocess.run(["cuobjdump", "-elf", obj], capture_output=True, text=True).stdout
want = sys.argv[1]
mn = sys.argv[2] if len(sys.argv) > 2 else None
funcs = sorted(set(re.findall(r"(_ZN3vr5\w+)", names)))
for fn in funcs:
    if want not in fn:
        continue
    sass = subprocess.run(["cuobjdump", "-sass", "-fun", fn, obj], capture_output=True, text=True).stdout
    ins = [(int(m.group(1), 16), m.group(2).strip()) for m in re.finditer(r"/\*([0-9a-f]{4})\*/\s+(.*?);", sass)]
    addr = {a: i for i, (a, _) in enumerate(ins)}
    for i, (a, t) in enumerate(ins):
        m = re.search(r"BRA\s+(?:!?U?P\d,\s*)?(?:!?P\d,\s*)?0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) <= a and int(m.group(1), 16) in addr:
            j = addr[int(m.group(1), 16)]
            body = [x for _, x in ins[j:i + 1]]
            keys = [mn] if mn else ["TLD", "LDG.E.64", "LDG.E.U8"]
            if any(any(k in x for k in keys) for x in body) and len(body) < 400:
                cnt = {}
                for x in body:
                    op = x.split()[1] if x.startswith("@") else x.split()[0]
                    cnt[op.split(".")[0]] = cnt.get(op.split(".")[0], 0) + 1
                print(fn[:60], "loop 0x%x..0x%x: %d instr" % (int(m.group(1), 16), a, len(body)), {k: v for k, v in sorted(cnt.items(), key=lambda kv: -kv[1])[:8]}, "LDL/STL:", sum("LDL" in x or "STL" in x for x in body))
