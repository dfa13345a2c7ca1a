"""Checkpoint round trip and the progress watchdog (CPU)."""
import time

import numpy as np

from psb200.models import LogRegLayout
from psb200.utils import checkpoint
from psb200.utils.watchdog import Watchdog


def test_checkpoint_roundtrip_uses_reference_keys(tmp_path):
    lay = LogRegLayout(64, 5)
    W = lay.random(3)
    path = checkpoint.save(str(tmp_path), 7, W, lay, clocks=[7, 7], rows_ingested=123)
    lines = open(path).read().splitlines()
    assert lines[0] == "key;value" and len(lines) == 1 + lay.num_params
    k, v = lines[1 + lay.key_of(3, 2)].split(";")
    assert int(k) == lay.key_of(3, 2) and abs(float(v) - W[3, 2]) < 1e-7
    W2, meta = checkpoint.load(str(tmp_path), lay)
    assert np.array_equal(W2, W) and meta["clock"] == 7 and meta["rows_ingested"] == 123


def test_watchdog_trips_on_stall_and_on_status():
    msgs = []
    wd = Watchdog(lambda: 5, lambda: 0, timeout_s=0.3, on_stall=msgs.append, poll_s=0.05).start()
    time.sleep(0.8)
    wd.stop()
    assert wd.tripped and "no progress" in msgs[0]
    msgs.clear()
    wd = Watchdog(lambda: int(time.time() * 100), lambda: 2, timeout_s=5, on_stall=msgs.append, poll_s=0.05).start()
    time.sleep(0.3)
    wd.stop()
    assert wd.tripped and "status word 2" in msgs[0]


def test_watchdog_quiet_while_progressing():
    wd = Watchdog(lambda: int(time.time() * 100), lambda: 0, timeout_s=0.2, poll_s=0.05).start()
    time.sleep(0.6)
    wd.stop()
    assert wd.tripped is None
