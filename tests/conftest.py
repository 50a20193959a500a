import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


# The oracle is the checker of every GPU parity test: on the GPU box (`-m gpu`, no /root/reference) it is first
# re-validated against the committed golden vectors, so these CPU tests are selected by BOTH marker expressions.
CHECKER_FILES = ("test_oracle_golden.py", "test_parity_util.py")


@pytest.hookimpl(tryfirst=True)
def pytest_collection_modifyitems(config, items):
    if "gpu" in (config.option.markexpr or "") and "not gpu" not in (config.option.markexpr or ""):
        for item in items:
            if os.path.basename(str(item.fspath)) in CHECKER_FILES:
                item.add_marker(pytest.mark.gpu)
                item.user_properties.append(("checker_on_gpu_box", True))
    try:
        import torch
        has_cuda = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_cuda = False
    if has_cuda:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords and not any(k == "checker_on_gpu_box" for k, _ in item.user_properties):
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
