"""The documents name files and tests; every one of them must exist (COVERAGE.md is read line by line)."""
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
