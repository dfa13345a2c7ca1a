"""psb_worker_step_static — the worker step for hosts that may not allocate device memory.

Some embedders of libpsb200.so cannot call cudaMalloc once they are running: the GPU's memory is held by
another framework's pool, or the loop is latency-critical and must not take the device-wide
synchronisation an allocation implies.  For them the library carries a fixed scratch array in its own
device image.  This test drives that entry point through the plain C ABI (ctypes, raw device pointers),
twice in a row on the same array, and checks both results against psb_worker_step with a caller-provided
scratch buffer.  It also checks that an oversized request is refused instead of overrunning the array.

The file name sorts last on purpose: if this addition ever breaks, a `pytest -x` run has already
exercised every other test.
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu


def _setup(native, F, cap, n, W, seed):
    dev = torch.device("cuda")
    rng = np.random.default_rng(seed)
    X = (rng.random((W, cap, F)) < 0.3) * rng.standard_normal((W, cap, F))
    X /= np.maximum(np.linalg.norm(X, axis=2, keepdims=True), 1e-6)
    y = rng.integers(1, 6, size=(W, cap)).astype(np.int32)
    win = np.tile(np.array([[n % cap, n]], dtype=np.int32), (W, 1))
    return (torch.from_numpy(X.astype(np.float16)).to(dev), torch.from_numpy(y).to(dev),
            torch.from_numpy(win).to(dev))


@pytest.mark.parametrize("impl", [0, 1])
def test_static_scratch_step_matches_regular_step_twice(impl):
    from psb200.ops import native
    L = native.lib()
    L.psb_static_scratch_bytes.restype = C.c_size_t
    vp = C.c_void_p
    L.psb_worker_step_static.argtypes = [C.POINTER(native.PsbSolver), C.POINTER(native.PsbComm), C.c_int, C.c_int,
                                         vp, vp, vp, vp, vp, vp, vp, vp]
    L.psb_worker_step_static.restype = C.c_int
    dev = torch.device("cuda")
    F, cap, n, W = 1024, 256, 200, 3
    nv = (F + 1) * 8
    assert L.psb_static_scratch_bytes() >= W * native.scratch_floats(F) * 4
    sv = native.make_solver(F, 6, cap, "lbfgs2", 2, True, True, 0.5)
    comm = native.comm_none()
    stream = torch.cuda.current_stream().cuda_stream
    scratch = torch.zeros((W, native.scratch_floats(F)), dtype=torch.float32, device=dev)
    for call in range(2):                                   # two calls in a row share the built-in array
        X, y, win = _setup(native, F, cap, n, W, seed=10 + call)
        w_in = torch.from_numpy((np.random.default_rng(call).standard_normal(nv) * 0.02).astype(np.float32)).to(dev)
        w_in.view(F + 1, 8)[:, 6:] = 0
        out = {}
        for name in ("regular", "static"):
            delta = torch.zeros((W, nv), dtype=torch.float32, device=dev)
            loss = torch.zeros(W, dtype=torch.float32, device=dev)
            info = torch.zeros((W, 8), dtype=torch.int32, device=dev)
            if name == "regular":
                rc = L.psb_worker_step(C.byref(sv), C.byref(comm), W, impl, X.data_ptr(), y.data_ptr(), win.data_ptr(),
                                       w_in.data_ptr(), delta.data_ptr(), loss.data_ptr(), info.data_ptr(),
                                       scratch.data_ptr(), stream)
            else:
                rc = L.psb_worker_step_static(C.byref(sv), C.byref(comm), W, impl, X.data_ptr(), y.data_ptr(),
                                              win.data_ptr(), w_in.data_ptr(), delta.data_ptr(), loss.data_ptr(),
                                              info.data_ptr(), stream)
            assert rc == 0, L.psb_last_error().decode()
            torch.cuda.synchronize()
            out[name] = (delta.cpu().numpy(), loss.cpu().numpy(), info.cpu().numpy())
        d0, l0, i0 = out["regular"]
        d1, l1, i1 = out["static"]
        assert np.isfinite(l1).all() and np.abs(d0).max() > 0
        assert np.array_equal(i0[:, :4], i1[:, :4])                             # evals, iterations, status, rows
        assert np.abs(l0 - l1).max() < 2e-3 * max(1.0, np.abs(l0).max())        # same tolerance as test_gpu_kernels
        assert np.abs(d0 - d1).max() < 2e-3 * max(1e-3, np.abs(d0).max())


def test_static_scratch_refuses_oversized_request():
    from psb200.ops import native
    L = native.lib()
    vp = C.c_void_p
    L.psb_worker_step_static.argtypes = [C.POINTER(native.PsbSolver), C.POINTER(native.PsbComm), C.c_int, C.c_int,
                                         vp, vp, vp, vp, vp, vp, vp, vp]
    L.psb_worker_step_static.restype = C.c_int
    dev = torch.device("cuda")
    F, cap, W = 1024, 64, 5                                 # one worker more than the array can serve
    sv = native.make_solver(F, 6, cap, "sgd", 2, True, True, 0.5)
    comm = native.comm_none()
    dummy = torch.zeros(16, dtype=torch.float32, device=dev)
    rc = L.psb_worker_step_static(C.byref(sv), C.byref(comm), W, 0, dummy.data_ptr(), dummy.data_ptr(),
                                  dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(),
                                  dummy.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == -3                                         # PSB_E_LIMIT
    msg = L.psb_last_error().decode()
    assert "built-in array" in msg and "5 workers" in msg
