import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vulkan-pathtracer_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import orc
    orc.build()
    return orc.lib()


@pytest.fixture(scope="session")
def ref_glsl():
    """The reference's brdf.glsl / rand.glsl compiled as C++ (oracle/_ref); built from /root/reference
    when present, prebuilt on the GPU box, otherwise the dependent tests are skipped."""
    from oracle import orc
    orc.build()
    lib = orc.ref_lib()
    if lib is None:
        pytest.skip("oracle/_ref/libref_glsl.so not available (no /root/reference and no prebuilt copy)")
    return lib
