"""CPU: the C-ABI library builds/loads, exports every symbol the header declares, and fails loudly without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from nexus_b200 import build, _lib
    build.build()                      # no-op when up to date; nvcc cross-compiles sm_100a without a GPU
    return _lib.load()


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "nexus_flat.h")).read()
    return sorted(set(re.findall(r"NXF_API\s+[\w\s\*]+?\b(nxf_\w+)\s*\(", hdr)))


def test_header_declares_expected_entry_points():
    syms = declared_symbols()
    for must in ["nxf_create", "nxf_add", "nxf_search", "nxf_merge_topk", "nxf_pool_normalize", "nxf_destroy",
                 "nxf_last_error", "nxf_ntotal", "nxf_reset", "nxf_reconstruct"]:
        assert must in syms


def test_library_exports_every_declared_symbol(lib):
    from nexus_b200 import _lib
    syms = declared_symbols()
    assert sorted(_lib.EXPORTS) == syms, "python binding and header disagree"
    for s in syms:
        assert hasattr(lib, s), f"{s} not exported"
    assert b"sm_100a" in lib.nxf_version()


def test_binary_contains_blackwell_kernels():
    """The product path is hand-written sm_100a: tcgen05 MMA, TMEM loads and TMA must be in the SASS."""
    import shutil
    import subprocess
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not on PATH")
    obj = os.path.join(ROOT, "nexus_b200", "_build", "search_tc.o")
    if not os.path.exists(obj):
        pytest.skip("object file not present (prebuilt .so only)")
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass and "LDTM" in sass and "UTMALDG" in sass
    assert "UTCHMMA.2CTA" in sass


def test_no_gpu_means_loud_failure_not_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("this check is for the CPU-only container")
    from nexus_b200 import FlatIndex, NexusFlatError
    import nexus_b200
    with pytest.raises(NexusFlatError) as e:
        FlatIndex(32)
    assert e.value.code in (-2, -4)
    with pytest.raises(RuntimeError):
        nexus_b200.index(corpus_embeddings=np.zeros((4, 8), np.float32))
    with pytest.raises(RuntimeError):
        nexus_b200.pool_and_normalize(torch.zeros(1, 2, 8), torch.ones(1, 2), "cls")


def test_argument_validation_without_gpu(lib):
    h = C.c_void_p()
    assert lib.nxf_create(0, 0, 0, 0, C.byref(h)) == -1 and b"dim" in lib.nxf_last_error()
    assert lib.nxf_create(8, 7, 0, 0, C.byref(h)) == -1 and b"metric" in lib.nxf_last_error()
    assert lib.nxf_create(8, 0, 9, 0, C.byref(h)) == -1 and b"dtype" in lib.nxf_last_error()
    assert lib.nxf_search(None, None, 1, 0, 0, 0, 1, None, None, 0, 0, None) == -1
    assert lib.nxf_ntotal(None) == -1
    assert lib.nxf_pool_normalize(None, 0, None, 1, 1, 1, 0, 1, 1e-12, None, 0, 0, None) == -1
    assert lib.nxf_pool_normalize(C.c_void_p(8), 0, None, 1, 1, 1, 9, 1, 1e-12, C.c_void_p(8), 0, 0, None) == -1
    assert b"not implemented" in lib.nxf_last_error()          # base.py:378-379


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under nexus_b200/ may reference it."""
    pkg = os.path.join(ROOT, "nexus_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f
                assert "/root/reference" not in src, f


@pytest.mark.parametrize("k,s,n", [(128, 135, 34375), (128, 2149, 34375), (1256, 135, 34375), (1256, 2149, 34375), (26, 148, 78125),
                                   (632, 2516, 78125), (16, 17, 4297), (128, 269, 4297), (2048, 64, 1024), (8, 1, 1024)])
def test_sample_rank_bounds_the_failure_probability(lib, k, s, n):
    """nxf_sample_rank (host arithmetic): the number of global top-k rows inside an evenly spread sample of s/n of the corpus is
    hypergeometric; the search fails to find k candidates above the sample's m-th score only if that count reaches m.  The
    chosen m keeps that probability below e^-30 (checked with the exact hypergeometric and Poisson tails), is minimal for
    the Chernoff bound it is derived from, and never exceeds k."""
    import math
    from scipy import stats
    m = lib.nxf_sample_rank(k, s, n)
    assert 1 <= m <= k
    lam = k * s / n
    if m < k:      # m == k is the guaranteed bound (the sample's k-th best), no probability involved
        chern = lambda mm: mm * math.log(mm / lam) - mm + lam
        assert chern(m) >= 30.0
        assert m - 1 <= lam or chern(m - 1) < 30.0
        assert stats.poisson.logsf(m - 1, lam) <= -30.0
        rows = 256
        assert stats.hypergeom.logsf(m - 1, n * rows, k, s * rows) <= -30.0
    assert lib.nxf_sample_rank(0, s, n) < 0 and lib.nxf_sample_rank(k, n + 1, n) < 0


def test_header_is_plain_c_and_a_c_program_links(tmp_path):
    """The boundary is a C ABI: include/nexus_flat.h must compile as C (no C++ in the signatures) and a C program must link
    against the shared library and run its host-only entry points without a GPU."""
    import shutil
    import subprocess
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("gcc not found")
    from nexus_b200 import _lib as L
    inc = os.path.join(ROOT, "include")
    src = tmp_path / "probe.c"
    src.write_text(
        '#include <stdio.h>\n#include "nexus_flat.h"\n'
        "int main(void) {\n"
        "  nxf_index* idx = 0;\n"
        '  printf("%s\\n", nxf_version());\n'
        "  int m = nxf_sample_rank(128, 1000, 16000);\n"
        "  int rc = nxf_search_subsets(idx, 0, 1, NXF_F32, 0, 0, 0, 0, 0, 0, 10, 0, 0, 0, 0);\n"   # NULL index: NXF_EINVAL
        '  printf("%d %d %d\\n", m, rc, (int)nxf_ntotal(idx));\n'
        "  return (m >= 1 && m <= 128 && rc == NXF_EINVAL) ? 0 : 1;\n}\n")
    subprocess.run([gcc, "-std=c99", "-Wall", "-Werror", "-pedantic", "-fsyntax-only", "-I", inc, str(src)], check=True)
    exe = tmp_path / "probe"
    subprocess.run([gcc, "-std=c99", "-I", inc, str(src), "-o", str(exe), L.LIB_PATH, "-Wl,-rpath," + os.path.dirname(L.LIB_PATH)], check=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, (r.stdout, r.stderr)
    assert r.stdout.startswith("nexus_flat ")
