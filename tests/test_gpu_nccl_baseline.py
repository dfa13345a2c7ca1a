"""The NCCL-baseline engine computes the same sequential-consistency trajectory as the fused engine."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from psb200.config import PSConfig
from psb200.data import ReviewGenerator

pytestmark = pytest.mark.gpu


def test_baseline_engine_matches_fused_engine_single_gpu():
    from psb200.engine.device_engine import DeviceEngine
    from psb200.parallel.nccl_baseline import NcclBaselineEngine
    F, cap, W = 256, 128, 3
    cfg = PSConfig(num_workers=W, num_features=F, max_buffer_size=cap, min_buffer_size=16)
    gen = ReviewGenerator(F, 5, vocab=3000, mean_len=120.0)
    b = gen.train_split(W * cap)
    rows = torch.from_numpy(b.to_dense(np.float16)).view(W, cap, F).cuda()
    labs = torch.from_numpy(b.labels).view(W, cap).cuda()
    out = []
    for cls in (DeviceEngine, NcclBaselineEngine):
        e = cls(cfg, W, impl=1)
        e.fill(rows, labs)
        for _ in range(3):
            e.step()
        out.append((e.weights().cpu().numpy().copy(), e.read_loss().numpy().copy()))
        e.close()
    assert np.abs(out[0][0] - out[1][0]).max() < 1e-4 * max(1.0, np.abs(out[0][0]).max())
    assert np.abs(out[0][1] - out[1][1]).max() < 1e-4
