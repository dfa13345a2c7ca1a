"""Documents, header, library and bindings must agree: the files and tests the documents name exist (COVERAGE.md is
read line by line), the header's functions are exported and listed, the ctypes mirrors match the C structs."""
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "parameter-server-architecture-on-apache-kafka_b200"
DOCS = ["COVERAGE.md", "INTEGRATION.md", "README.md", "DESIGN.md", "profiles/README.md", "evaluation/README.md",
        "kubernetes/README.md"]
OF_THE_REFERENCE = {"evaluation/python-ground-truth-algorithm.ipynb", "kubernetes/kafka.yaml"}     # cited, not ours


def test_every_file_and_test_named_in_the_documents_exists():
    tests = set()
    for f in glob.glob(os.path.join(ROOT, "tests", "*.py")):
        tests.update(re.findall(r"^def (test_\w+)", open(f).read(), re.M))
    stale = set()
    for doc in DOCS:
        text = open(os.path.join(ROOT, doc)).read()
        for tok in re.findall(r"`([^`\n]+)`", text):
            stale.update((doc, t) for t in re.findall(r"::(test_\w+)", tok) if t not in tests)
            for path, more in re.findall(r"(?:^|[\s(,])((?:P|csrc|tests|tools|evaluation|profiles|kubernetes|dev|mockData|"
                                         r"include)/[\w./\-]+)([*{]?)", " " + tok):
                path = path.rstrip(".,:").split("::")[0]
                real = os.path.join(ROOT, path.replace("P/", PKG + "/", 1) if path.startswith("P/") else path)
                found = bool(glob.glob(real + "*")) if more else os.path.exists(real)      # `csrc/ps_eval*.cu`: a prefix
                if path not in OF_THE_REFERENCE and not found:
                    stale.add((doc, path))
        if doc == "profiles/README.md":
            for name in re.findall(r"`((?:r[12]_|sass/)[\w./\-]+)`", text):
                if not os.path.exists(os.path.join(ROOT, "profiles", name)):
                    stale.add((doc, name))
    assert not stale, sorted(stale)


def test_header_library_and_integration_table_agree():
    """Every function include/psb200.h declares is exported by the built library and named in INTEGRATION.md's
    table; the library exports no undeclared psb_ function (kernels' host stubs aside)."""
    import subprocess

    import pytest

    from psb200.ops import native
    if not native.available():
        pytest.skip("libpsb200.so not built")
    h = open(os.path.join(ROOT, "include", "psb200.h")).read()
    h = re.sub(r"//[^\n]*", "", re.sub(r"/\*.*?\*/", "", h, flags=re.S))
    declared = set(re.findall(r"\b(psb_\w+)\s*\(", h))
    lib = native.lib()
    assert not [n for n in declared if not hasattr(lib, n)]
    out = subprocess.run(["nm", "-D", "--defined-only", lib._name], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (psb_\w+)", out))
    assert not [n for n in exported if n not in declared and not n.endswith("_kernel")]
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    named = set(re.findall(r"psb_\w+", doc))
    for stem, parts in re.findall(r"(psb_\w+?_)(\w+(?:/\w+)+)", doc):            # psb_producer_start/poll/release/stop
        named.update(stem + p for p in parts.split("/"))
    for stem, parts in re.findall(r"(psb_\w+)`((?: / `_\w+`)+)", doc):             # `psb_comm_version_words` / `_rank_clocks`
        base = stem.rsplit("_", 2)[0] if stem.count("_") > 2 else stem
        for suffix in re.findall(r"`(_\w+)`", parts):
            named.update({base + suffix, "_".join(stem.split("_")[:2]) + suffix, "_".join(stem.split("_")[:3]) + suffix})
    wild = [w[:-1] for w in re.findall(r"psb_\w+_\*", doc)]                        # psb_tracker_*
    missing = [n for n in declared if n not in named and not any(n.startswith(w) for w in wild)]
    assert not missing, sorted(missing)


def test_ctypes_mirrors_match_the_header_field_by_field(tmp_path):
    """P/ops/native.py restates the structs of include/psb200.h for ctypes; a C program compiled from the header
    prints the offset and size of every field, which must be what ctypes lays out."""
    import ctypes
    import shutil
    import subprocess

    import pytest

    from psb200.ops import native
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no C compiler")
    mirrors = {"PsbSolver": native.PsbSolver, "PsbComm": native.PsbComm, "PsbClockRecord": native.PsbClockRecord,
               "PsbSparseComm": native.PsbSparseComm, "PsbSlabComm": native.PsbSlabComm}
    lines = ['#include <stdio.h>', f'#include "{os.path.join(ROOT, "include", "psb200.h")}"', "int main(void) {"]
    for name, cls in mirrors.items():
        lines.append(f'  printf("{name} . %zu %zu\\n", (size_t)0, sizeof({name}));')
        for field, _ in cls._fields_:
            lines.append(f'  printf("{name} {field} %zu %zu\\n", offsetof({name}, {field}), sizeof((({name}*)0)->{field}));')
    lines += ["  return 0;", "}"]
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines) + "\n")
    exe = tmp_path / "layout"
    r = subprocess.run([cc, "-std=c11", str(src), "-o", str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr                        # a field the mirror has and the header lacks fails here
    out = subprocess.run([str(exe)], capture_output=True, text=True).stdout.split("\n")
    seen = 0
    for ln in filter(None, out):
        name, field, off, size = ln.split()
        cls = mirrors[name]
        if field == ".":
            assert ctypes.sizeof(cls) == int(size), (name, ctypes.sizeof(cls), size)
        else:
            f = getattr(cls, field)
            assert (f.offset, f.size) == (int(off), int(size)), (name, field, f.offset, f.size, off, size)
        seen += 1
    assert seen == sum(len(c._fields_) + 1 for c in mirrors.values())
    # and the header has no field the mirror lacks: the sizes above would differ unless it were padding
    h = open(os.path.join(ROOT, "include", "psb200.h")).read()
    assert native.MAX_RANKS == int(re.search(r"#define\s+PSB_MAX_RANKS\s+(\d+)", h).group(1))
