"""What of bench.py's contract with the driver can be checked without a GPU: the reference arm (the reference is
Java / Kafka / Spark and cannot be installed here: one JSON line saying so, exit 0, from rank 0 only), the defaults
(N = 1, at least 3 warm-up steps), and that a GPU-less box gets an error instead of a CPU fallback."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BENCH = os.path.join(ROOT, "bench.py")


def _run(args, **env):
    return subprocess.run([sys.executable, BENCH] + args, capture_output=True, text=True, timeout=300,
                          env=dict(os.environ, **env), cwd=ROOT)


def test_reference_arm_prints_one_line_saying_why_it_cannot_run():
    r = _run(["--impl", "reference", "--gpus", "1", "--steps", "5", "--warmup", "3"])
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert r.returncode == 0 and len(lines) == 1
    out = json.loads(lines[0])
    assert out["impl"] == "reference" and isinstance(out["unavailable"], str) and "\n" not in out["unavailable"]
    # launched like the other arm for N > 1: every rank exits 0, only rank 0 prints
    r = _run(["--impl", "reference", "--gpus", "8"], RANK="3", WORLD_SIZE="8", LOCAL_RANK="3")
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_defaults_and_no_cpu_fallback():
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_module", BENCH)
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    argv, sys.argv = sys.argv, ["bench.py"]
    try:
        a = bench.parse()
    finally:
        sys.argv = argv
    assert a.gpus == 1 and a.warmup >= 3 and a.steps >= 1 and a.impl == "ours" and a.transport == "p2p"
    assert a.consistency_model == 0 and a.solver == "lbfgs2" and (a.workers_per_gpu, a.buffer, a.features) == (296, 1024, 1024)
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if not has_gpu:                                   # our arm on a box without a GPU: an error, never a silent CPU run
        r = _run(["--steps", "3", "--warmup", "3"])
        assert r.returncode != 0 and r.stdout.strip() == ""


def test_graft_entry_build_compiles_and_imports():
    """__graft_entry__.build(): every source compiled for sm_100a in-tree (incremental: up-to-date objects are kept),
    package importable, library loadable — without a GPU."""
    import shutil

    import pytest
    if shutil.which("nvcc") is None and not os.path.exists("/usr/local/cuda/bin/nvcc"):
        pytest.skip("no nvcc")
    sys.path.insert(0, ROOT)
    import __graft_entry__ as entry
    entry.build()
    from psb200.ops import native
    assert native.available() and native.lib().psb_version() >= 100
    assert callable(entry.smoke)
    flags = open(os.path.join(ROOT, "build.sh")).read()
    assert "arch=compute_100a,code=sm_100a" in flags and "-lineinfo" in flags
