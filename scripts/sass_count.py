"""Static SASS profile of one kernel in a library: opcode histogram of the whole function and of its hottest
backward-branch loop (largest loop body).  usage: sass_count.py <lib.so> <substring of the mangled name>"""
import collections, re, subprocess, sys
lib, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
for f in funcs[1:]:
    name = f.split("\n", 1)[0].strip()
    if pat not in name or "_param_" in name:
        continue
    ins = []
    for line in f.split("\n"):
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m:
            ins.append((int(m.group(1), 16), m.group(2)))
    print(name, "instructions:", len(ins))
    loops = []
    for addr, t in ins:
        m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?(?:`\(\.L_x_\d+\)|0x([0-9a-f]+))", t)
        m2 = re.search(r"BRA.*0x([0-9a-f]+)", t)
        if m2:
            tgt = int(m2.group(1), 16)
            if tgt < addr:
                loops.append((addr - tgt, tgt, addr))
    loops.sort(reverse=True)
    for span, a, b in loops[:3]:
        body = [t for ad, t in ins if a <= ad <= b]
        h = collections.Counter(re.sub(r"^@!?U?P\d\s+", "", t).split()[0].split(".")[0] for t in body)
        print(f"  loop 0x{a:x}..0x{b:x}: {len(body)} instr;", ", ".join(f"{k} {v}" for k, v in h.most_common(14)))
    if "--dump" in sys.argv and loops:
        span, a, b = loops[0]
        for ad, t in ins:
            if a <= ad <= b:
                print(f"    {ad:05x}  {t}")
