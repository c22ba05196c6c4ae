"""The C-ABI library loads, exports every function include/nodalproj.h declares, and refuses to run without a GPU
(there is no CPU compute path in the product).  No compute calls here."""
import ctypes as C
import os
import re
import subprocess

import pytest

from incflo_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions(header="nodalproj.h", prefix="np_"):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(" + prefix + r"[a-z_0-9]+)\s*\(", src)))


def product_lib():
    if not os.path.exists(capi.LIB_PATH):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "incflo_b200", "csrc")])
    return capi.load()


def test_header_and_binding_agree():
    assert declared_functions() == sorted(capi.SYMBOLS)
    assert declared_functions("macproj.h", "mp_") == sorted(capi.MAC_SYMBOLS)


def test_library_exports_every_declared_symbol():
    L = product_lib()
    missing = [s for s in declared_functions() + declared_functions("macproj.h", "mp_") if not hasattr(L, s)]
    assert not missing, missing


def test_no_link_time_nccl_dependency():
    out = subprocess.run(["ldd", capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "nccl" not in out          # bound at run time so the host process keeps its own NCCL


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    L = product_lib()
    g = capi.np_geom()
    for d in range(3):
        g.ncell[d] = 8; g.prob_lo[d] = 0.0; g.prob_hi[d] = 1.0; g.periodic[d] = 1
    bx = (capi.np_box * 1)()
    for d in range(3):
        bx[0].lo[d], bx[0].hi[d] = 0, 7
    owner = (C.c_int * 1)(0)
    h = C.c_void_p()
    rc = L.np_create(C.byref(h), C.byref(g), 1, bx, owner, 0, 1, None, None)
    assert rc == capi.NP_ERR_NO_DEVICE
    assert b"no CPU fallback" in L.np_strerror(rc)


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(ImportError):
        capi.load(str(tmp_path / "libnodalproj.so"))
