import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

os.environ.setdefault("ATTACKFL_QUIET", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200 box)")
    config.addinivalue_line("markers", "slow: long-running statistical / end-to-end test")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference's ``src.Utils`` / ``src.Model`` (oracle)."""
    from baseline import ref_loader
    if not ref_loader.ensure_installed():
        pytest.skip("reference not installed under baseline/_ref")

    class R:
        utils = ref_loader.load_utils()
        model = ref_loader.load_model()
        loader = ref_loader
    return R
