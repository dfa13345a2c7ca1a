"""The C++ host runtime under ThreadSanitizer and AddressSanitizer + UBSan (`make tsan`, `make asan`): the producer
thread with tokenising helpers against a consumer that holds several batches, the packers and the tokenizer on buffers
of exactly the documented size, awkward text, the buffer policy's error paths.  Host only."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CANNOT_START = ("unexpected memory mapping", "Shadow memory range interleaves", "cannot find -ltsan", "cannot find -lasan",
                "cannot find -lubsan", "ReserveShadowMemoryRange failed")


@pytest.mark.parametrize("target,done", [("tsan", "tsan driver: 3000 rows consumed, 3000 sent"),
                                         ("asan", "buffer policy and tracker checked")])
def test_host_runtime_is_clean_under(target, done):
    if shutil.which("make") is None or shutil.which("g++") is None:
        pytest.skip("needs make and g++")
    r = subprocess.run(["make", target], cwd=ROOT, capture_output=True, text=True, timeout=900)
    out = r.stdout + r.stderr
    if r.returncode != 0 and any(s in out for s in CANNOT_START):
        pytest.skip("the sanitizer runtime cannot start in this environment")
    assert r.returncode == 0 and done in out, out[-3000:]
    assert "WARNING: ThreadSanitizer" not in out and "ERROR: AddressSanitizer" not in out and "runtime error" not in out
