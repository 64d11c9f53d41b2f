import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


@pytest.fixture(scope="session")
def oracle():
    """Strongest CPU oracle available (compiled reference, else the CPU port). TEST INFRASTRUCTURE."""
    import oracle_api
    try:
        return oracle_api.load()
    except oracle_api.OracleError as e:
        pytest.skip(str(e))


@pytest.fixture(scope="session")
def ref_oracle():
    """The compiled, unmodified reference; tests needing it are skipped where it cannot exist (GPU box from a clean checkout)."""
    import oracle_api
    if not oracle_api.available("reference"):
        pytest.skip("oracle/_ref/libspgref.so not built (needs /root/reference)")
    return oracle_api.load("reference")


@pytest.fixture(scope="session")
def ctx():
    from spg_b200 import api
    c = api.Context(0)
    yield c
    c.close()
