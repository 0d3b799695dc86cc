import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import __graft_entry__ as graft  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def rt():
    """The package (vulkan-raytracing_b200/) under its importable name, native libraries built."""
    mod = graft.load_package()
    need = [mod.HOST_LIB_PATH, mod.RT_LIB_PATH, os.path.join(ROOT, "oracle", "liboracle.so"),
            os.path.join(ROOT, "oracle", "liboracle64.so"),
            os.path.join(ROOT, "tests", "emu", "libemu.so")]
    if not all(os.path.exists(p) for p in need):
        graft.build()
    return mod


@pytest.fixture(scope="session")
def ob(rt):
    import oracle_binding
    return oracle_binding


@pytest.fixture(scope="session")
def emu(rt):
    import emu_binding
    return emu_binding


@pytest.fixture(scope="session")
def gpu_rt(rt):
    if rt.device_count() < 1:
        pytest.fail("gpu-marked test started without a CUDA device: there is no CPU fallback")
    return rt
