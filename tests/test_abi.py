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


def test_struct_layouts_match_the_headers(tmp_path):
    """The ctypes mirrors of the POD structs (np_geom, np_box, np_fab, np_options, np_stats, np_apply_args, mp_options) have the
    size and the field offsets a C compiler gives the declarations in include/*.h."""
    import ctypes as C
    import shutil
    import subprocess
    from incflo_b200 import capi
    if not shutil.which("gcc"):
        pytest.skip("no C compiler")
    structs = {"np_geom": capi.np_geom, "np_box": capi.np_box, "np_fab": capi.np_fab, "np_options": capi.np_options,
               "np_stats": capi.np_stats, "np_apply_args": capi.np_apply_args, "mp_options": capi.mp_options}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "nodalproj.h"', '#include "macproj.h"', 'int main(void) {']
    for name, cls in structs.items():
        lines.append(f'  printf("{name} size %zu\\n", sizeof({name}));')
        for fname, _ in cls._fields_:
            lines.append(f'  printf("{name} {fname} %zu\\n", offsetof({name}, {fname}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split("\n")
    for ln in out:
        if not ln.strip():
            continue
        name, field, val = ln.split()
        cls = structs[name]
        want = C.sizeof(cls) if field == "size" else getattr(cls, field).offset
        assert int(val) == want, (name, field, int(val), want)
