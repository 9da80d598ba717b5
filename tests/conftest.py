import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


@pytest.fixture(scope="session")
def O():
    """The CPU oracle (test infrastructure)."""
    from oracle import oracle as orc
    orc.build()
    return orc


@pytest.fixture(scope="session")
def rt():
    import raytracingproject_b200 as pkg
    return pkg


@pytest.fixture(scope="session")
def scene(O):
    return O.scene_rtiow()


@pytest.fixture(scope="session")
def ctx(rt):
    """A device context on cuda:0 -- through the C ABI; fails loudly without a GPU."""
    c = rt.Context(0)
    yield c
    c.close()


@pytest.fixture
def ctx_scene(ctx, scene, rt):
    """the session's context with the final scene (reference src/main.cpp:11-53) uploaded -- again for every test that asks
    for it, because other tests upload other worlds into the same context"""
    s, m = scene
    ctx.upload_scene(s.view(rt.SPHERE), m.view(rt.MATERIAL))
    return ctx


def psnr8(a, b):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)
