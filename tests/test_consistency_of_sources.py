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


def test_ctypes_signatures_match_the_header_prototypes():
    """Every prototype of include/psb200.h against the argtypes / restype P/ops/native.py declares: the same number
    of parameters, pointers where the header has pointers, 4-byte and 8-byte integers and floats where it has them
    (an int passed for a size_t leaves the upper half of the register undefined)."""
    import ctypes as C

    import pytest

    from psb200.ops import native
    if not native.available():
        pytest.skip("libpsb200.so not built")
    lib = native.lib()
    h = open(os.path.join(ROOT, "include", "psb200.h")).read()
    h = re.sub(r"//[^\n]*", "", re.sub(r"/\*.*?\*/", "", h, flags=re.S))
    scalars = {"int": "i4", "unsigned": "i4", "unsigned int": "i4", "uint32_t": "i4", "int32_t": "i4", "size_t": "i8",
               "long long": "i8", "unsigned long long": "i8", "uint64_t": "i8", "int64_t": "i8", "float": "f4",
               "double": "f8"}

    def of_c(param):
        param = param.strip()
        if param in ("void", ""):
            return None
        if "*" in param or "[" in param or "cudaStream_t" in param:
            return "ptr"
        return scalars[re.sub(r"\b\w+$", "", param).replace("const", "").strip()]

    def of_py(a):
        if a in (C.c_void_p, C.c_char_p) or (isinstance(a, type) and issubclass(a, C._Pointer)):
            return "ptr"
        return {C.c_int: "i4", C.c_uint: "i4", C.c_size_t: "i8", C.c_longlong: "i8", C.c_ulonglong: "i8",
                C.c_float: "f4", C.c_double: "f8"}[a]

    protos = re.findall(r"([\w\s\*]+?)\b(psb_\w+)\s*\(([^;{]*?)\)\s*;", h)
    assert len(protos) >= 69
    for ret, name, params in protos:
        want = [k for k in map(of_c, params.split(",")) if k]
        f = getattr(lib, name)
        assert f.argtypes is not None or not want, f"{name}: no argtypes declared"
        assert [of_py(a) for a in (f.argtypes or [])] == want, name
        ret = ret.replace("const", "").strip()
        if "*" in ret:
            assert f.restype in (C.c_void_p, C.c_char_p), name
        elif ret != "void" and scalars[ret] == "i8":
            assert f.restype in (C.c_size_t, C.c_longlong, C.c_ulonglong), name     # default restype is a 4-byte int
