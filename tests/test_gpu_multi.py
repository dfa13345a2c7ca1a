"""Real multi-GPU run of the peer-memory protocol (one process per GPU, cudaIpc-mapped blocks)."""
import os
import subprocess
import sys

import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("model,impl", [(0, 1), (0, 0), (-1, 1), (2, 1)])
def test_two_ranks_over_peer_memory(model, impl):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = 29500 + (os.getpid() + 7 * model + impl) % 400
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tools", "mp_check.py"), "--model", str(model), "--impl", str(impl)],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "-> OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("path", ["direct", "slab"])
def test_two_ranks_sharded_sparse(path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = 29900 + (os.getpid() + len(path)) % 90
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tools", "mp_check_sparse.py"), "--path", path],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "-> OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("model", [2, -1])
def test_per_worker_gates_hold_across_real_gpus(model):
    """Bounded delay k / eventual between WORKERS on different GPUs: in-kernel gates on the replicated clock table,
    a straggler on the last rank (tools/mp_worker_clocks.py asserts max lag == k, or that nobody waits)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    n = min(torch.cuda.device_count(), 8)
    port = 29700 + (os.getpid() + 13 * (model + 2)) % 150
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n),
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tools", "mp_worker_clocks.py"), "--model", str(model)],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "-> OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("model", [0, -1, 4])
def test_node_app_on_two_gpus_writes_server_and_worker_rows(tmp_path, model):
    """The fused entry point under torchrun for every consistency model: rank 0 scores the global weights each clock
    (gathered from the owners' shards in the relaxed models) and every rank writes its worker rows."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import numpy as np
    from psb200.utils.logs import parse_log
    port = 29300 + (os.getpid() + 17 * (model + 2)) % 150
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), "-m", "psb200.apps.node", "-training", "synthetic:4000", "-test",
           "synthetic:800", "-p", "0", "--workers_per_gpu", "2", "-max", "256", "-l", "-c", str(model), "--max_clock", "6",
           "--clocks_per_launch", "2" if model != 0 else "1"]
    r = subprocess.run(cmd, env=env, cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    s = parse_log(str(tmp_path / "logs-server.csv"), "server")
    w0 = parse_log(str(tmp_path / "logs-worker.csv"), "worker")
    w1 = parse_log(str(tmp_path / "logs-worker-1.csv"), "worker")
    assert len(s) == (6 if model == 0 else 3) and s[-1, 2] == 5 and s[-1, 4] > 0.2 and np.all(np.isfinite(s[:, 4]))
    assert len(w0) == len(w1) == 12 and set(w0[:, 1]) == {0, 1} and set(w1[:, 1]) == {2, 3}
    assert np.all(np.isfinite(w0[:, 3])) and np.all(np.isfinite(w1[:, 3]))


@pytest.mark.parametrize("model", [2, -1])
def test_sharded_sparse_path_bounded_delay_and_eventual_with_a_straggler(model):
    """Bounded delay on the 2^24-style sharded model (the reference's gate does not depend on the model size,
    ServerProcessor.java:126-131): fast ranks advance at the straggler's pace after k free steps; eventual ranks do not."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = 29100 + (os.getpid() + 11 * (model + 2)) % 150
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tools", "mp_sparse_bounded.py"), "--model", str(model)],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "-> OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
