import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: takes more than a few seconds")


@pytest.fixture(scope="session")
def golden():
    from tests.util import load_golden
    return load_golden


def pytest_terminal_summary(terminalreporter):
    """say when a fit passed on the extended-precision criterion rather than against the reference-pinned FP64 checker"""
    from tests import util
    for u in util.SECOND_CRITERION_USES:
        terminalreporter.write_line("[parity] extended-precision criterion used: N,K,M,A=%s  CUDA-checker %.2e  CUDA-ext %.2e  "
                                    "checker-ext %.2e" % (u["shape"], u["e_gw"], u["e_ge"], u["e_we"]))
