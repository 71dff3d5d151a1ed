import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def synth_dir(tmp_path_factory):
    return str(tmp_path_factory.mktemp("synth"))


def pytest_terminal_summary(terminalreporter):
    """Exemption counters of every parity run (helpers.report), so they show in the tail of a -q log."""
    try:
        from helpers import PARITY_REPORT
    except Exception:
        return
    if PARITY_REPORT:
        terminalreporter.write_line("parity exemption counters (0 = bit-exact sets / every ray within 1e-5):")
        for line in PARITY_REPORT:
            terminalreporter.write_line("  " + line)
