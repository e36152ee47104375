import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def lib():
    from raytracer_project_b200 import _lib
    return _lib.load()


@pytest.fixture(scope="session")
def gpu_ctx():
    from raytracer_project_b200 import api
    ctx = api.Context(0)
    ctx.set_tuning("build_check", 1)  # every device-built tree is structurally checked (csrc/sah_build.cu validate_tree)
    yield ctx
    ctx.close()
