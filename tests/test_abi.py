"""The C-ABI library loads on a CPU-only box, exports every symbol the header declares,
and refuses to compute without a GPU (no fallback).  CPU only."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from sequenza_b200 import _lib, build as build_mod

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "sequenza_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sqz_[a-z0-9_]+)\s*\(", src)))


def test_library_is_built_in_tree():
    build_mod.build()
    assert os.path.exists(_lib.SO_PATH)
    assert os.path.dirname(_lib.SO_PATH) == os.path.join(ROOT, "sequenza_b200")


def test_every_declared_symbol_is_exported():
    handle = C.CDLL(_lib.SO_PATH)
    names = declared_symbols()
    assert len(names) >= 17
    for n in names:
        assert hasattr(handle, n), "header declares %s but the library lacks it" % n
    assert set(names) == set(_lib.SIGNATURES), "ctypes table and header disagree"


def test_version_and_error_strings():
    L = _lib.lib()
    assert b"sm_100a" in L.sqz_version()
    assert isinstance(L.sqz_last_error(), bytes)
    assert L.sqz_baf_grid_workspace_bytes(1000, 1, 61, 91) % 256 == 0
    assert L.sqz_baf_grid_workspace_bytes(100000, 1, 601, 901) > 100000 * 32


def test_sm100a_only():
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    out = subprocess.run([cuobjdump, "-lelf", _lib.SO_PATH], capture_output=True, text=True)
    archs = set(re.findall(r"sm_\d+a?", out.stdout))
    assert archs == {"sm_100a"}, archs


def test_no_gpu_means_error_not_fallback():
    L = _lib.lib()
    if L.sqz_device_count() > 0:
        pytest.skip("a GPU is visible; the refusal path is for CPU-only hosts")
    import sequenza_b200 as S
    with pytest.raises(RuntimeError, match="no CUDA device"):
        S.baf_bayes([0.4], [1.0], 0.5, 2.0, 1.0)
    with pytest.raises(RuntimeError, match="no CUDA device"):
        S.baf_model_fit(Bf=[0.4], depth_ratio=[1.0], avg_depth_ratio=1.0)
    with pytest.raises(RuntimeError, match="no CUDA device"):
        S.mufreq_bayes([0.4], [1.0], 0.5, 2.0, 1.0)
    out = np.zeros(4)
    x = np.ones(1)
    rc = L.sqz_baf_grid_dev(1, None, 1, *[x.ctypes.data] * 6, x.ctypes.data, 1,
                            x.ctypes.data, 1, 0, 1, x.ctypes.data, 0, 7, 2, None, None, 0, 0,
                            0, None, 0, None, out.ctypes.data)
    assert rc == -3


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "sequenza_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".c", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.lower(), "%s mentions the oracle" % f


def test_header_is_plain_c_and_links(tmp_path):
    """include/sequenza_b200.h compiles as C99, every declared entry point links, and the
    program runs: on a CPU-only host the compute call reports SQZ_ERR_NOGPU (exit 0)."""
    import shutil
    import subprocess
    gcc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else shutil.which("gcc")
    if not gcc:
        pytest.skip("no C compiler")
    src = os.path.join(ROOT, "examples", "abi_check.c")
    text = open(src).read()
    for name in declared_symbols():
        assert "(fn_t)%s" % name in text, "examples/abi_check.c does not reference %s" % name
    exe = str(tmp_path / "abi_check")
    libdir = os.path.dirname(_lib.SO_PATH)
    subprocess.check_call([gcc, "-std=c99", "-Wall", "-Werror", "-pedantic",
                           "-I" + os.path.join(ROOT, "include"), src, "-L" + libdir,
                           "-lsequenza_b200", "-Wl,-rpath," + libdir, "-o", exe])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "sm_100a" in out.stdout
