"""profiles/sass/MNEMONICS.md from `cuobjdump -sass` of the built library (runs without a GPU).
Usage: python tools/sass_summary.py"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "parameter-server-architecture-on-apache-kafka_b200", "ops", "libpsb200.so")
OUT = os.path.join(ROOT, "profiles", "sass", "MNEMONICS.md")
TENSOR = ("UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMAPF", "UTMACCTL", "UTCBAR", "UTCATOMSWS", "SYNCS", "UTCCP")
NOTABLE = ("LDG", "STG", "LDS", "STS", "LD.E", "ST.E", "RED", "ATOM", "FFMA", "HFMA2", "MUFU", "ELECT", "R2UR", "MEMBAR", "REDUX",
           "LDL", "STL", "SHFL", "POPC", "CCTL")

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
kernels = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = kernels.setdefault(m.group(1), collections.Counter())
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m and cur is not None:
        cur[m.group(1)] += 1

with open(OUT, "w") as f:
    f.write("# SASS mnemonic summary per kernel (cuobjdump -sass libpsb200.so, sm_100a; tools/sass_summary.py)\n\n"
            "Tensor-path mnemonics: UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = TMA tensor load, UTCBAR = "
            "tcgen05.commit, UTCATOMSWS = tcgen05.alloc / dealloc, SYNCS.* = mbarrier ops.  `.STRONG.SYS` on LDG / RED / "
            "ATOMG = system-scope accesses to peer memory.\n\n")
    for name, c in kernels.items():
        total = sum(c.values())
        tens = {k: v for k, v in sorted(c.items()) if k.startswith(TENSOR)}
        note = {k: v for k, v in sorted(c.items()) if k.startswith(NOTABLE) and not k.startswith(TENSOR)}
        f.write(f"## `{name}`\n\n{total} instructions.  tensor/TMA/mbarrier: {tens if tens else 'none'}\n\n"
                f"other notable: {note}\n\n")
print("wrote", OUT, "for", len(kernels), "kernels")
