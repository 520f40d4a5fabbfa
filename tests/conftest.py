import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_terminal_summary(terminalreporter):
    """Worst observed error of every tolerance comparison, relative to max(|oracle|, scale)."""
    from tests.util import WORST
    if WORST:
        terminalreporter.write_line("parity: worst |got - oracle| / max(|oracle|, scale) per comparison (bar 1e-12)")
        for what, w in sorted(WORST.items()):
            terminalreporter.write_line(f"  {what:44s} {w:.3e}")


@pytest.fixture(scope="session")
def eng():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but no CUDA device is visible (there is no CPU fallback)")
    from k8s_rocm_metrics_dashboard_b200 import Engine
    e = Engine(0)
    yield e
    e.close()


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")
