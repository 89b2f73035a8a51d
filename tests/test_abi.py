"""CPU: liblpslam_cuda.so loads without a GPU and exports every symbol include/lpslam_cuda.h declares; argument checks
that need no device answer with the documented status codes; compute calls fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "lpslam_cuda.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lpslam_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    from lpslam_b200 import _lib, knn, match  # noqa: F401  (register the late-bound entry points)
    L = _lib.lib()
    names = declared_symbols()
    assert len(names) >= 35
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(line.split()[-1] for line in out.splitlines() if line.strip())
    for n in names:
        assert n in exported, f"{n} declared in include/lpslam_cuda.h but not exported"
        getattr(L, n)
    # every symbol the Python mirror binds is declared in the header
    for n in _lib.SIGNATURES:
        assert n in names, f"{n} bound by the mirror but missing from the header"


def test_no_torch_or_oracle_in_product():
    """the product package never imports the oracle, and the shared library links neither torch nor the oracle"""
    from lpslam_b200 import _lib
    for f in os.listdir(os.path.join(ROOT, "lpslam_b200")):
        if f.endswith(".py"):
            src = open(os.path.join(ROOT, "lpslam_b200", f)).read()
            assert not re.search(r"^\s*(from|import)\s+(oracle|\.\.oracle)", src, flags=re.M), f"{f} imports the oracle"
    out = subprocess.run(["ldd", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "torch" not in out and "orb_oracle" not in out and "orb_ref" not in out


def test_abi_version_and_errors_without_device():
    from lpslam_b200 import _lib
    L = _lib.lib()
    assert L.lpslam_cuda_abi_version() == 1
    h = C.c_void_p()
    p = _lib.OrbParams(2000, 4000, 1.2, 8, 20, 7, None, 0)
    # argument errors are reported before any device call
    assert L.lpslam_orb_create(None, C.byref(p), 0, 1) == _lib.ERR_ARG
    bad = _lib.OrbParams(2000, 4000, 1.2, 99, 20, 7, None, 0)
    assert L.lpslam_orb_create(C.byref(h), C.byref(bad), 0, 1) == _lib.ERR_ARG
    assert b"num_levels" in L.lpslam_cuda_last_error()
    rects = (C.c_float * 4)(0.5, 0.2, 0.1, 0.6)  # x_min > x_max: orb_params.cc:16-26 throws
    badr = _lib.OrbParams(2000, 4000, 1.2, 8, 20, 7, rects, 1)
    assert L.lpslam_orb_create(C.byref(h), C.byref(badr), 0, 1) == _lib.ERR_ARG
    import torch
    if not torch.cuda.is_available():
        # no CPU fallback: creation fails with a CUDA error instead of silently computing on the host
        rc = L.lpslam_orb_create(C.byref(h), C.byref(p), 0, 1)
        assert rc == _lib.ERR_CUDA and h.value is None
        with pytest.raises(_lib.LpslamCudaError):
            from lpslam_b200 import orb
            orb.orb_extractor()


def test_python_mirror_param_errors():
    from lpslam_b200 import orb
    with pytest.raises(RuntimeError):
        orb.orb_params(mask_rects=[[0.5, 0.2, 0.1, 0.6]])
    with pytest.raises(RuntimeError):
        orb.orb_params(mask_rects=[[0.1, 0.2, 0.7, 0.6]])
    with pytest.raises(RuntimeError):
        orb.orb_params(mask_rects=[[0.1, 0.2, 0.7]])
    # the five exception vectors of the reference's own test/openvslam/feature/orb_params.cc:60-166
    for bad in ([0.28, 0.59, 0.1], [0.28, 0.28, 0.1, 0.2], [0.25, 0.21, 0.1, 0.2], [0.28, 0.59, 0.8, 0.8], [0.28, 0.59, 0.7, 0.2]):
        with pytest.raises(RuntimeError):
            orb.orb_params(mask_rects=[[0.2, 0.5, 0.3, 0.8], bad])
    ok = orb.orb_params(max_num_keypts=3000, ini_max_num_keypts=5000, scale_factor=1.3, num_levels=12, ini_fast_thr=25, min_fast_thr=9,
                        mask_rects=[[0.2, 0.5, 0.3, 0.8], [0.28, 0.59, 0.1, 0.2]])  # :24-58
    assert (ok.max_num_keypts_, ok.ini_max_num_keypts_, ok.num_levels_, ok.ini_fast_thr_, ok.min_fast_thr) == (3000, 5000, 12, 25, 9)
    assert len(ok.mask_rects_) == 2 and np.float32(ok.mask_rects_[1][1]) == np.float32(0.59)
    p = orb.orb_params()
    assert (p.max_num_keypts_, p.num_levels_, p.ini_fast_thr_, p.min_fast_thr) == (2000, 8, 20, 7)
    assert np.float32(p.scale_factor_) == np.float32(1.2)
