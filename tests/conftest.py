import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(autouse=True)
def _knobs_follow_the_environment():
    """Tests that set RP_* knobs (monkeypatch + reload_env) must not leak them: the library reads its
    knob table once, so after every test that loaded it the table is re-read from the restored
    environment (this fixture is set up first, hence torn down after monkeypatch)."""
    yield
    mod = sys.modules.get("instancesegmentation_jittor_b200._lib")
    if mod is not None and getattr(mod, "_lib", None) is not None:
        mod.reload_env()
