import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

SCENES = os.path.join(ROOT, "scenes") + "/"
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built_lib():
    """libmcpt.so, (re)built if sources are newer; nvcc cross-compiles without a GPU."""
    from montecarlopathtracing_b200 import build
    return build.build()


@pytest.fixture(scope="session")
def oracle(built_lib):
    import oracle_py
    oracle_py.lib()
    return oracle_py


@pytest.fixture(scope="session")
def scenes(built_lib):
    import montecarlopathtracing_b200 as M
    return {name: M.HostScene(SCENES, name) for name in ("cornell-box", "veach-mis")}


@pytest.fixture(scope="session")
def gpu_ctx(built_lib):
    import montecarlopathtracing_b200 as M
    ctx = M.Context(0)  # raises if there is no B200: no CPU fallback
    yield ctx
    ctx.close()
