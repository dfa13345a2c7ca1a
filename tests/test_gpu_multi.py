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
