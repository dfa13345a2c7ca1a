"""The built library is what profiles/sass/ shows: the two tensor kernels issue tcgen05 / TMA / TMEM instructions, the
fused kernels reach peer memory with system-scope accesses, and the committed listings are the listings of this build
(`make sass` regenerates them).  Needs cuobjdump (CUDA toolkit), no GPU."""
import collections
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "parameter-server-architecture-on-apache-kafka_b200", "ops", "libpsb200.so")
SASS = os.path.join(ROOT, "profiles", "sass")


@pytest.fixture(scope="module")
def kernels():
    if shutil.which("cuobjdump") is None or shutil.which("c++filt") is None or not os.path.exists(LIB):
        pytest.skip("needs cuobjdump, c++filt and the built library")
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    mangled = sorted(set(re.findall(r"Function : (\S+)", out)))
    plain = subprocess.run(["c++filt"] + mangled, capture_output=True, text=True).stdout.split("\n")
    names = {m: p.split("(")[0] for m, p in zip(mangled, plain)}
    found, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = found.setdefault(names[m.group(1)], [])
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if m and cur is not None:
            cur.append(m.group(1))
    return found


def test_tensor_kernels_issue_tcgen05_tma_and_tmem_instructions(kernels):
    for name in ("psb_worker_tc_kernel", "psb_eval_tc_kernel"):
        ops = collections.Counter(kernels[name])
        assert ops["UTCHMMA"] >= 1, name                                   # tcgen05.mma
        assert any(k.startswith("UTMALDG") for k in ops), name             # cp.async.bulk.tensor (TMA load)
        assert any(k.startswith("LDTM") for k in ops), name                # tcgen05.ld out of TMEM
        assert ops["UTCBAR"] >= 1, name                                    # tcgen05.commit -> mbarrier
        assert any(k.startswith("UTCATOMSWS") for k in ops), name          # tcgen05.alloc / dealloc
        assert any(k.startswith("SYNCS.PHASECHK") for k in ops), name      # mbarrier try_wait
        assert not any(k.startswith(("HMMA", "IMMA", "WGMMA")) for k in ops), name     # no legacy tensor path beside it


def test_fused_kernels_reach_peer_memory_with_system_scope_accesses(kernels):
    for name in ("psb_worker_tc_kernel", "psb_worker_simt_kernel", "psb_worker_csr_kernel", "psb_sparse_step_kernel"):
        ops = set(kernels[name])
        assert any(k.startswith("LDG") and k.endswith("STRONG.SYS") for k in ops), name        # ld.acquire / relaxed.sys
        assert any(k.startswith(("RED", "ATOM", "STG")) and "STRONG.SYS" in k for k in ops), name


def test_committed_listings_are_the_listings_of_this_build(kernels):
    files = sorted(f[:-5] for f in os.listdir(SASS) if f.endswith(".sass"))
    assert files == sorted(kernels), "profiles/sass/ lists other kernels than the library has: run `make sass`"
    for name, ops in kernels.items():
        listed = []
        for line in open(os.path.join(SASS, name + ".sass")):
            m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
            if m:
                listed.append(m.group(1))
        assert listed == ops, f"{name}: profiles/sass/{name}.sass is not this build's code: run `make sass`"
