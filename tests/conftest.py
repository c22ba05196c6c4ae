import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


import pytest  # noqa: E402


@pytest.hookimpl(tryfirst=True)
def pytest_collection_modifyitems(config, items):
    """A GPU-only parameter value (e.g. host fabs handed to the CUDA library) combined with the CPU-emulation backend is not
    a GPU test: take its gpu mark away, so that every test `-m gpu` selects drives CUDA, and skip it everywhere."""
    for it in items:
        cs = getattr(it, "callspec", None)
        if cs is not None and cs.params.get("backend") == "emu" and it.get_closest_marker("gpu") is not None:
            it.own_markers = [m for m in it.own_markers if m.name != "gpu"]
            it.add_marker(pytest.mark.skip(reason="GPU-only parameter on the emulation backend"))


@pytest.fixture(params=["emu", pytest.param("gpu", marks=pytest.mark.gpu)])
def backend(request):
    """Which build of the product sources a parity test drives through the C ABI.

    emu: TEST-ONLY host emulation of the CUDA execution model (tests/emu) -> runs on CPU, checks logic.
    gpu: the product, incflo_b200/libnodalproj.so, on cuda:0 with device-resident fabs (zero copy).
    """
    import nphelp
    if request.param == "emu":
        return dict(lib=nphelp.emu_lib(), like="numpy")
    import torch
    from incflo_b200 import capi
    assert torch.cuda.is_available(), "gpu-marked test needs a CUDA device"
    torch.cuda.set_device(0)
    return dict(lib=capi.load(), like="cuda", device="cuda:0")
