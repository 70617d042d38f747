"""profiles/r0N_sass_mnemonics.txt: which kernels of the built library use cp.async (LDGSTS), the TMA engine (UBLKCP),
three-input min/max (FMNMX3) and vector float reductions (ATOM.E.ADD.F32x4 with no destination) - static counts."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "soillib_b200", "libsoilb200.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
KEYS = ("LDGSTS", "LDGDEPBAR", "UBLKCP", "UTMALDG", "SYNCS", "FMNMX3", "ATOM.E.ADD.F32x4", "RED", "LDG.E.128", "STG.E.128")
rows = []
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    name = f.split("\n", 1)[0].strip()
    short = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    short = re.sub(r"\(anonymous namespace\)::|sb::|^void ", "", short)
    short = re.sub(r"\(.*", "", short)
    ops, n = collections.Counter(), 0
    for line in f.split("\n"):
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Za-z0-9_.]+)", line)
        if m:
            n += 1
            for w in KEYS:
                if m.group(1).startswith(w):
                    ops[w] += 1
    if any(ops[w] for w in ("LDGSTS", "UBLKCP", "UTMALDG", "FMNMX3", "ATOM.E.ADD.F32x4")):
        rows.append((short, n, ops))
print("# SASS mnemonics per kernel of libsoilb200.so (cuobjdump -sass, sm_100a; static counts; scripts/sass_mnemonics.py)")
print("# LDGSTS / LDGDEPBAR = cp.async into shared memory; UBLKCP / SYNCS = cp.async.bulk (TMA engine) + mbarrier; UTMALDG = cp.async.bulk.tensor (2-D tensor-map loads);")
print("# FMNMX3 = three-input min/max; ATOM.E.ADD.F32x4 (destination RZ) = 16-byte float reductions of the flux records")
for short, n, ops in sorted(rows):
    print(f"{short[:64]:64s} {n:6d} instr  " + "  ".join(f"{k} {v}" for k, v in sorted(ops.items()) if v))
