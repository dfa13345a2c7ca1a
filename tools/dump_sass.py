"""Full `cuobjdump -sass` listing of EVERY kernel of libpsb200.so, one file per kernel under profiles/sass/
(runs without a GPU), plus the mnemonic summary (tools/sass_summary.py).  Usage: python tools/dump_sass.py"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "parameter-server-architecture-on-apache-kafka_b200", "ops", "libpsb200.so")
OUT = os.path.join(ROOT, "profiles", "sass")

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
chunks, cur, name = {}, None, None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        cur = chunks.setdefault(name, [f"// {m.group(1)}"])
        continue
    if cur is not None and line.strip() and not line.lstrip().startswith(".."):
        # drop the encoding column: the mnemonics are what the listing is for
        cur.append(re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", line.rstrip()))
os.makedirs(OUT, exist_ok=True)
for old in os.listdir(OUT):
    if old.endswith(".sass"):
        os.remove(os.path.join(OUT, old))
for name, lines in chunks.items():
    body = [ln for ln in lines if not re.match(r"^\s*/\* 0x[0-9a-f]{16} \*/\s*$", ln)]
    with open(os.path.join(OUT, name + ".sass"), "w") as f:
        f.write("\n".join(body) + "\n")
    print(f"{name}: {sum(1 for ln in body if re.match(r'^\s+/\*[0-9a-f]{4,}\*/', ln))} instructions")
subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sass_summary.py")], check=True)
