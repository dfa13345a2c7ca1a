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


def test_rows_file_streams_clean_under_asan(tmp_path):
    """The memory-mapped rows-file reader of the producer thread, into buffers of exactly the documented size."""
    exe = os.path.join(ROOT, "build", "asan_packers")
    if not os.path.exists(exe):
        pytest.skip("make asan has not built the driver (previous test skipped)")
    from psb200.data import ReviewGenerator, write_rows_file
    from psb200.ops import native
    if not native.available():
        pytest.skip("libpsb200.so not built")
    batch = ReviewGenerator(num_features=64, vocab=400, mean_len=200).train_split(700)       # dense rows: long payloads
    path = str(tmp_path / "t.psbrows")
    write_rows_file(path, batch)
    r = subprocess.run([exe, path, "64", "700"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.count("700 rows") == 2, r.stdout + r.stderr
