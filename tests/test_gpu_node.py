"""The fused entry point end to end on one GPU: logs in the reference format, checkpoint and resume."""
import os
import subprocess
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")

from psb200.config import SERVER_LOG_HEADER, WORKER_LOG_HEADER
from psb200.utils.logs import parse_log

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))


def _node(cwd, *extra):
    cmd = [sys.executable, "-m", "psb200.apps.node", "-training", "synthetic:3000", "-test", "synthetic:800", "-p", "0",
           "--workers_per_gpu", "4", "-max", "256", "-l", *extra]
    return subprocess.run(cmd, env=ENV, cwd=cwd, capture_output=True, text=True, timeout=600)


@pytest.mark.parametrize("model", [0, -1])
def test_node_streams_learns_and_resumes(tmp_path, model):
    ck = str(tmp_path / "ck")
    r = _node(tmp_path, "--max_clock", "4", "-c", str(model), "--checkpoint_dir", ck, "--checkpoint_every", "2")
    assert r.returncode == 0, r.stderr[-3000:]
    s = open(tmp_path / "logs-server.csv").read().splitlines()
    w = open(tmp_path / "logs-worker.csv").read().splitlines()
    assert s[0] == SERVER_LOG_HEADER and w[0] == WORKER_LOG_HEADER
    srows, wrows = parse_log(s, "server"), parse_log(w, "worker")
    assert list(srows[:, 2]) == [0, 1, 2, 3] and len(wrows) == 16
    assert np.all(np.isfinite(wrows[:, 3])) and srows[-1, 4] > 0.2
    if model == 0:
        assert os.path.exists(os.path.join(ck, "weights-4.csv"))
        r = _node(tmp_path, "--max_clock", "6", "--checkpoint_dir", ck, "--resume")
        assert r.returncode == 0, r.stderr[-3000:]
        s2 = parse_log(str(tmp_path / "logs-server.csv"), "server")
        assert list(s2[:, 2]) == [4, 5]                    # resumed at the checkpointed clock
        assert s2[0, 4] > 0.2                              # ... with the checkpointed weights, not zeros


def test_node_streams_a_csv_file_through_the_producer_thread(tmp_path):
    from psb200.data import ReviewGenerator, write_csv
    gen = ReviewGenerator(num_features=256, vocab=2000, mean_len=60.0)
    write_csv(str(tmp_path / "train.csv"), gen.train_split(1500))
    write_csv(str(tmp_path / "test.csv"), gen.test_split(400))
    cmd = [sys.executable, "-m", "psb200.apps.node", "-training", "train.csv", "-test", "test.csv", "-p", "0",
           "--num_features", "256", "--workers_per_gpu", "2", "-max", "256", "-l"]
    r = subprocess.run(cmd, env=ENV, cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    s = open(tmp_path / "logs-server.csv").read().splitlines()
    assert "CSV Producer sent all records!" in s
    srows = parse_log(s, "server")
    wrows = parse_log(str(tmp_path / "logs-worker.csv"), "worker")
    assert len(srows) >= 2 and wrows[:, 6].max() == 750          # each of the 2 workers saw half of the 1500 rows
    assert srows[-1, 4] > srows[0, 4] - 0.05


def test_worker_metrics_and_dynamic_buffer(tmp_path):
    """--worker_metrics fills the worker rows' F1/accuracy; arrival timestamps size the device buffer."""
    r = _node(tmp_path, "--max_clock", "3", "--worker_metrics")
    assert r.returncode == 0, r.stderr[-3000:]
    w = parse_log(str(tmp_path / "logs-worker.csv"), "worker")
    assert np.all((w[:, 4] >= 0) & (w[:, 4] <= 1)) and np.all((w[:, 5] >= 0) & (w[:, 5] <= 1))
    from psb200.config import PSConfig
    from psb200.engine.device_engine import DeviceEngine
    eng = DeviceEngine(PSConfig(num_workers=2, max_buffer_size=1024, min_buffer_size=128), 2, impl=0, stage_rows=4096)
    rows = torch.zeros((2000, 1024), dtype=torch.float16).pin_memory()
    labs = torch.ones(2000, dtype=torch.int32).pin_memory()
    workers = (np.arange(2000) % 2).astype(np.int32)
    eng.ingest(rows, labs, workers, np.arange(2000) // 2 * 30.0)        # 2000 events/min per worker -> 0.3*2000 = 600
    torch.cuda.synchronize()
    assert eng.win_host[:, 1].tolist() == [600, 600] and eng.win.cpu()[:, 1].tolist() == [600, 600]
    eng.ingest(rows[:400], labs[:400], workers[:400], 30000.0 + np.arange(400) // 2 * 1000.0)   # 60/min -> shrinks
    assert all(128 <= c < 600 for c in eng.win_host[:, 1].tolist())
    eng.close()


def test_server_app_runner_with_cuda_backend(tmp_path):
    """The reference-style entry point with the sm_100a kernels doing the math (in-process, sequential)."""
    cmd = [sys.executable, "-m", "psb200.apps.server_app_runner", "-training", "synthetic:1200", "-test",
           "synthetic:500", "-p", "0", "--num_workers", "2", "--in_process", "--max_clock", "4", "-l", "--backend", "cuda",
           "-max", "256"]
    r = subprocess.run(cmd, env=ENV, cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    s = parse_log(str(tmp_path / "logs-server.csv"), "server")
    w = parse_log(str(tmp_path / "logs-worker.csv"), "worker")
    assert len(s) == 4 and len(w) >= 6 and np.all(np.isfinite(w[:, 3])) and s[-1, 4] > 0.2


def test_sparse_and_dense_ingestion_fill_identical_rings():
    from psb200.config import PSConfig
    from psb200.data import random_csr
    from psb200.engine.device_engine import DeviceEngine
    cfg = PSConfig(num_workers=3, max_buffer_size=128, min_buffer_size=16, num_features=256)
    b = random_csr(200, 256, 30, seed=8)
    workers = (np.arange(200) % 3).astype(np.int32)
    engs = [DeviceEngine(cfg, 3, impl=0, stage_rows=256) for _ in range(2)]
    engs[0].ingest(torch.from_numpy(b.to_dense(np.float16)).pin_memory(),
                   torch.from_numpy(b.labels.astype(np.int32)).pin_memory(), workers)
    engs[1].ingest_sparse(b, workers)
    torch.cuda.synchronize()
    assert torch.equal(engs[0].X, engs[1].X) and torch.equal(engs[0].labels, engs[1].labels)
    assert torch.equal(engs[0].win.cpu(), engs[1].win.cpu())
    for e in engs:
        e.close()


@pytest.mark.parametrize("fmt", ["dense", "palette", "csr"])
def test_batch_larger_than_a_ring_leaves_the_last_rows(fmt):
    """More rows for one worker in ONE batch than its ring holds: the ring must end up with the last `cap` rows
    (the store kernels write the rows of a batch in parallel; superseded rows are dropped on the host)."""
    from psb200.config import PSConfig
    from psb200.data import random_csr
    from psb200.engine.device_engine import DeviceEngine
    from psb200.ops import native
    cap, F, n = 128, 256, 300
    cfg = PSConfig(num_workers=1, max_buffer_size=cap, min_buffer_size=16, num_features=F)
    b = random_csr(n, F, 30, seed=3)
    dense = b.to_dense(np.float16)
    eng = DeviceEngine(cfg, 1, impl=0, stage_rows=512)
    workers = np.zeros(n, dtype=np.int32)
    if fmt == "csr":
        eng.ingest_sparse(b, workers)
    elif fmt == "palette":
        eng.ingest(native.PackedRows(n, F, palette=True).pack_dense(dense, b.labels), None, workers)
    else:
        eng.ingest(torch.from_numpy(dense).pin_memory(), torch.from_numpy(b.labels.astype(np.int32)).pin_memory(), workers)
    torch.cuda.synchronize()
    X, y = eng.X[0].cpu().numpy(), eng.labels[0].cpu().numpy()
    for i in range(n - cap, n):                               # row i sits in slot i % cap
        assert np.array_equal(X[i % cap], dense[i]) and y[i % cap] == b.labels[i]
    assert eng.win_host[0].tolist() == [n % cap, cap]
    eng.close()


def test_checkpoint_restores_rings_buffer_statistics_and_relaxed_shards(tmp_path):
    """state_dict / load_state_dict: a resumed engine continues exactly like the one that kept running."""
    from psb200.config import PSConfig
    from psb200.data import ReviewGenerator
    from psb200.engine.device_engine import DeviceEngine
    from psb200.utils import checkpoint
    F, cap, W = 256, 128, 2
    for model, gran in ((0, "rank"), (-1, "worker")):
        cfg = PSConfig(num_workers=W, num_features=F, consistency_model=model, max_buffer_size=cap, min_buffer_size=16)
        b = ReviewGenerator(F, 5, vocab=3000, mean_len=120.0).train_split(W * 100)
        rows = torch.from_numpy(b.to_dense(np.float16)).pin_memory()
        labs = torch.from_numpy(b.labels.astype(np.int32)).pin_memory()
        a = DeviceEngine(cfg, W, impl=1, async_granularity=gran, stage_rows=256)
        a.ingest(rows, labs, (np.arange(W * 100) % W).astype(np.int32), np.arange(W * 100) * 40.0)
        for _ in range(3):
            a.step()
        torch.cuda.synchronize()
        checkpoint.save_rank_state(str(tmp_path), 0, a.state_dict())
        w_ck = a.weights().cpu().numpy().copy()
        c = DeviceEngine(cfg, W, impl=1, async_granularity=gran, stage_rows=256)
        c.load_state_dict(checkpoint.load_rank_state(str(tmp_path), 0), weights=w_ck)
        assert c.clock == 3 and c.rows_ingested == W * 100 and torch.equal(c.X, a.X) and torch.equal(c.win.cpu(), a.win.cpu())
        assert [x.ring_total() for x in c.buffers] == [100, 100]
        assert np.array_equal(c.buffers[0].get_state(), a.buffers[0].get_state())
        for e in (a, c):
            e.step()
        la, lc = a.read_loss().numpy().copy(), c.read_loss().numpy().copy()
        tol = 2e-5 if model == 0 else 2e-3            # the workers' REDs land in any order (float sums)
        assert np.allclose(la, lc, rtol=tol, atol=tol)
        wa = a.weights().cpu().numpy()
        assert np.abs(wa - c.weights().cpu().numpy()).max() < tol * max(1.0, np.abs(wa).max())
        a.close(), c.close()


def test_all_workers_local_models_scored_in_one_launch():
    """evaluate_workers (psb_eval_models) == evaluate() on each worker's local weights, SIMT and tcgen05."""
    from psb200.config import PSConfig
    from psb200.data import ReviewGenerator
    from psb200.engine.device_engine import DeviceEngine
    W, cap = 5, 256
    cfg = PSConfig(num_workers=W, max_buffer_size=cap)
    gen = ReviewGenerator(mean_len=300.0)
    tr, te = gen.train_split(W * cap), gen.test_split(700)
    for impl in (0, 1):
        e = DeviceEngine(cfg, W, impl=impl)
        e.fill(torch.from_numpy(tr.to_dense(np.float16)).view(W, cap, -1).cuda(), torch.from_numpy(tr.labels).view(W, cap).cuda())
        e.load_test(te)
        e.step(), e.step()
        f1, acc = e.evaluate_workers()
        for j in range(W):
            f1j, accj = e.evaluate(e.worker_local_weights(j).contiguous())
            assert abs(f1j - f1[j]) < 1e-6 and abs(accj - acc[j]) < 1e-6
        assert len(set(np.round(f1, 6))) > 1                  # the local models differ
        e.close()


@pytest.mark.parametrize("model,worker_device,extra", [(0, 0, []), (-1, 0, []), (3, 0, []), (0, 1, []), (-1, 1, []),
                                                       (0, 0, ["--free_running"])])
def test_two_runner_processes_train_through_device_peer_memory(tmp_path, model, worker_device, extra):
    """WorkerAppRunner and ServerAppRunner as two processes (run.sh with BACKEND=cuda): the rendezvous hands over two
    CUDA IPC handles, rows travel through the INPUT_DATA topic in device memory, weights and gradients through the
    server's peer-memory block inside the worker kernels; both logs come out in the reference format."""
    import socket
    if worker_device >= torch.cuda.device_count():
        pytest.skip("needs a second GPU: the worker app on another device than the server's block")
    with socket.socket() as so:
        so.bind(("127.0.0.1", 0))
        port = so.getsockname()[1]
    addr = f"127.0.0.1:{port}"
    wcmd = [sys.executable, "-m", "psb200.apps.worker_app_runner", "-l", "-test", "synthetic:600", "--backend", "cuda",
            "-max", "256", "--rendezvous", addr, "--device", str(worker_device)]
    scmd = [sys.executable, "-m", "psb200.apps.server_app_runner", "-l", "-p", "0", "-training", "synthetic:3000", "-test",
            "synthetic:600", "--backend", "cuda", "--num_workers", "3", "-c", str(model), "--max_clock", "5",
            "--rendezvous", addr, "--clocks_per_launch", "1", *extra]
    wdir, sdir = tmp_path / "w", tmp_path / "s"
    wdir.mkdir(), sdir.mkdir()
    wp = subprocess.Popen(wcmd, env=ENV, cwd=wdir, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    try:
        sr = subprocess.run(scmd, env=ENV, cwd=sdir, capture_output=True, text=True, timeout=600)
        wout, werr = wp.communicate(timeout=120)
    finally:
        if wp.poll() is None:
            wp.kill()
    assert sr.returncode == 0, sr.stdout[-1500:] + sr.stderr[-3000:]
    assert wp.returncode == 0, wout[-1500:] + werr[-3000:]
    s = open(sdir / "logs-server.csv").read().splitlines()
    w = open(wdir / "logs-worker.csv").read().splitlines()
    assert s[0] == SERVER_LOG_HEADER and w[0] == WORKER_LOG_HEADER
    srows, wrows = parse_log(s, "server"), parse_log(w, "worker")
    assert len(wrows) == 15 and sorted(set(wrows[:, 1])) == [0, 1, 2] and sorted(set(wrows[:, 2])) == [0, 1, 2, 3, 4]
    assert np.all(np.isfinite(wrows[:, 3])) and np.all((wrows[:, 4] > 0) & (wrows[:, 4] <= 1))    # local models scored
    if extra:                                              # free running: the server logs the clocks it gets to see
        assert 1 <= len(srows) <= 5 and srows[-1, 2] == 4 and np.all(np.diff(srows[:, 2]) > 0)
    else:
        assert list(srows[:, 2]) == [0, 1, 2, 3, 4]        # a server row for every clock of worker 0 (Q6)
    assert srows[-1, 4] > 0.2 and np.all(srows[:, 1] == -1) and np.all(srows[:, 3] == -1)


def test_runner_cli_streams_a_csv_file_over_the_p2p_transport(tmp_path):
    """ServerAppRunner --in_process --backend cuda on a real CSV: C++ producer thread -> topic slots -> rings."""
    from psb200.data import ReviewGenerator, write_csv
    gen = ReviewGenerator(num_features=256, vocab=2000, mean_len=60.0)
    write_csv(str(tmp_path / "train.csv"), gen.train_split(1200))
    write_csv(str(tmp_path / "test.csv"), gen.test_split(300))
    cmd = [sys.executable, "-m", "psb200.apps.server_app_runner", "-training", "train.csv", "-test", "test.csv", "-p", "0",
           "--num_features", "256", "--num_workers", "2", "--in_process", "-l", "--backend", "cuda", "-max", "256"]
    r = subprocess.run(cmd, env=ENV, cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    s = open(tmp_path / "logs-server.csv").read().splitlines()
    assert "CSV Producer sent all records!" in s
    w = parse_log(str(tmp_path / "logs-worker.csv"), "worker")
    assert w[:, 6].max() == 600 and np.all(np.isfinite(w[:, 3])) and parse_log(s, "server")[-1, 4] > 0.15


def test_store_into_spare_ring_slots_while_a_step_runs_gives_the_serialized_result():
    """ring_spare_rows + ingest_commit(overlap=True): batch i+1 is stored while step i trains; every step must see
    exactly the windows (and produce exactly the losses / weights) of the engine that stores between two steps."""
    from psb200.config import PSConfig
    from psb200.data import ReviewGenerator
    from psb200.engine.device_engine import DeviceEngine
    F, cap, W, r, steps = 256, 256, 4, 64, 9
    cfg = PSConfig(num_workers=W, num_features=F, max_buffer_size=cap, min_buffer_size=16)
    gen = ReviewGenerator(F, 5, vocab=3000, mean_len=120.0)
    batches = []
    for i in range(steps + 1):
        b = gen.train_split(W * r, seed=100 + i)
        batches.append((torch.from_numpy(b.to_dense(np.float16)).pin_memory(),
                        torch.from_numpy(b.labels.astype(np.int32)).pin_memory()))
    workers = np.repeat(np.arange(W, dtype=np.int32), r)
    out = []
    for spare in (0, r):
        e = DeviceEngine(cfg, W, impl=1, stage_rows=W * r, ring_spare_rows=spare)
        assert e.cap == cap + (128 if spare else 0)
        e.ingest_begin(*batches[0], workers)
        e.ingest_commit()
        losses = []
        for i in range(steps):
            e.ingest_begin(*batches[i + 1], workers)
            e.step()
            t = e.read_loss_begin()
            e.ingest_commit(overlap=bool(spare))              # spare == 0: orders itself behind the step
            losses.append(e.read_loss_end(t).numpy().copy())
        torch.cuda.synchronize()
        out.append((np.asarray(losses), e.weights().cpu().numpy().copy(), e.win_host.clone()))
        assert e.status() == 0
        e.close()
    # same rows, same windows; the rows sit in other ring slots, so only the summation order inside the GEMMs differs
    assert np.allclose(out[0][0], out[1][0], rtol=3e-4, atol=3e-5)
    assert np.abs(out[0][1] - out[1][1]).max() < 3e-4 * max(1e-2, np.abs(out[0][1]).max())
    assert out[0][2][:, 1].tolist() == [cap] * W               # the windows filled up and wrapped: (steps+1)*r > cap


def test_window_statistics_kept_at_ingestion_follow_growing_shrinking_and_wrapping_windows():
    """psb_window_stats_update: after every batch the 1/std table equals the statistics of the rows currently in
    each worker's window (unbiased std, 0 for constant columns), also when the dynamic buffer grows, shrinks and wraps."""
    from psb200.config import PSConfig
    from psb200.data import random_csr
    from psb200.engine.device_engine import DeviceEngine
    F, cap, W = 128, 256, 3
    cfg = PSConfig(num_workers=W, num_features=F, max_buffer_size=cap, min_buffer_size=32, buffer_size_coefficient=0.3)
    e = DeviceEngine(cfg, W, impl=0, stage_rows=1024, ring_spare_rows=128)
    assert e.stats_at_ingest and e.cap == cap + 128
    rng = np.random.default_rng(0)
    t = 0.0
    for step, (n, gap_ms) in enumerate([(90, 10.0), (128, 10.0), (300, 5.0), (60, 400.0), (45, 2000.0), (128, 20.0), (7, 20.0)]):
        b = random_csr(n * W, F, 20, seed=step)
        dense = b.to_dense(np.float16)
        dense[:, 5] = 0.25                                        # a constant non-zero column -> 1/std = 0
        workers = np.repeat(np.arange(W, dtype=np.int32), n)
        times = t + np.tile(np.arange(n) * gap_ms, W)
        t += n * gap_ms
        e.ingest(torch.from_numpy(dense).pin_memory(), torch.from_numpy(b.labels.astype(np.int32)).pin_memory(),
                 workers, times)
        torch.cuda.synchronize()
        X = e.X.cpu().numpy().astype(np.float64)
        inv = e._inv[e._win_i].cpu().numpy()
        for w in range(W):
            head, count = e.win_host[w].tolist()
            slots = [(head - count + i) % e.cap for i in range(count)]
            rows = X[w, slots]
            var = rows.var(axis=0, ddof=1) if count > 1 else np.zeros(F)
            ref = np.where(var > 1e-12, 1.0 / np.sqrt(np.maximum(var, 1e-300)), 0.0)
            assert np.allclose(inv[w], ref, rtol=1e-5, atol=1e-6), (step, w, count)
            assert inv[w, 5] == 0.0
    counts = e.win_host[:, 1].tolist()
    assert all(32 <= c <= cap for c in counts)
    e.close()
