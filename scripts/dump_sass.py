"""Regenerates profiles/sass/*.txt (one cuobjdump -sass listing per kernel of ops/libdvc_b200.so) and
profiles/sass/README.md (mnemonic counts that prove the Blackwell paths: UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld,
UTMALDG / UTMASTG = TMA, SYNCS = mbarrier).  No GPU needed:  python scripts/dump_sass.py"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributedvolunteercomputing_b200", "ops", "libdvc_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
os.makedirs(OUT, exist_ok=True)
txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
funcs = re.split(r"\n\s*Function : ", txt)[1:]
KEYS = ["UTCHMMA", "LDTM", "UTMALDG", "UTMASTG", "UTMAPF", "SYNCS", "FHFMA", "HMMA", "LDG", "STG", "LDS", "STS", "ATOM", "MEMBAR", "ACQBULK"]
rows = []
for f in funcs:
    name = f.split("\n", 1)[0].strip()
    pretty = demangle(name)
    short = re.sub(r"\(.*", "", pretty).replace("void ", "").replace("dvc::", "")
    fname = re.sub(r"[^A-Za-z0-9_]+", "_", short).strip("_")
    ops = collections.Counter()
    n = 0
    for line in f.splitlines():
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            ops[m.group(1)] += 1
            n += 1
    # drop the hex encodings (second line of every instruction, trailing /* 0x... */): mnemonics and operands stay
    f = "\n".join(re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", ln) for ln in f.splitlines() if not re.match(r"^\s*/\* 0x[0-9a-f]{16} \*/\s*$", ln))
    with open(os.path.join(OUT, f"sass_{fname}.txt"), "w") as fh:
        fh.write(f"// {pretty}\n// cuobjdump -sass ops/libdvc_b200.so (sm_100a), {n} instructions\n" + "Function : " + f)
    rows.append((short, fname, n, ops))
rows.sort()
with open(os.path.join(OUT, "README.md"), "w") as fh:
    fh.write("# SASS listings, one per kernel (`python scripts/dump_sass.py`)\n\n"
             "UTCHMMA = `tcgen05.mma`, LDTM = `tcgen05.ld`, UTMALDG / UTMASTG = TMA load / store, SYNCS = mbarrier, FHFMA = fp16 x fp16 + fp32 FMA.\n\n"
             "| kernel | file | instructions | " + " | ".join(KEYS) + " |\n|---|---|---|" + "---|" * len(KEYS) + "\n")
    for short, fname, n, ops in rows:
        fh.write(f"| `{short}` | `sass_{fname}.txt` | {n} | " + " | ".join(str(ops.get(k, 0) or "") for k in KEYS) + " |\n")
print(f"{len(rows)} kernels -> {OUT}")
