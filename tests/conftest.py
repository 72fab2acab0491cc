import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import util  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def rt():
    from _loader import load_build, load_package
    load_build().build()
    load_build().build_redo_variant()  # test build of tests/test_gpu_redo_build.py: never older than the sources
    return load_package()


@pytest.fixture(scope="session")
def ref():
    import ref_oracle
    if not ref_oracle.available():
        pytest.skip("oracle/_ref/libref_oracle.so not built (needs /root/reference in the build container)")
    return ref_oracle


@pytest.fixture(scope="session")
def bunny(rt):
    scene, mesh = util.bunny_scene(rt)
    return {"scene": scene, "mesh": mesh, "arrays": scene.arrays()}


@pytest.fixture(scope="session")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required for -m gpu tests (no CPU fallback exists)")
    return torch
