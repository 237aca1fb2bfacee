import os
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")
    config.addinivalue_line("markers", "ref: needs the compiled reference under oracle/_ref (built only where /root/reference exists)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure the product library and the oracle exist (no compute here)."""
    import __graft_entry__ as ge
    from vpic_b200.lib import B200_LIB
    from oracle import harness
    if not os.path.exists(B200_LIB) or not os.path.exists(harness.ORACLE_LIB):
        ge.build()
