import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built():
    """Build librtb200.so / the oracle if they are missing (no-op when the prebuilt files travelled)."""
    import subprocess
    pkg = os.path.join(ROOT, "ray-tracer-rust_b200")
    if not os.path.exists(os.path.join(pkg, "librtb200.so")):
        subprocess.check_call(["make", "-C", pkg, "-j4", "librtb200.so"])
    from oracle import oracle as O
    O.build()
    return True


@pytest.fixture(scope="session")
def big_bunny_inputs():
    import scenes as S
    v, c = S.big_bunny()
    return v, c, S.make_samples()


@pytest.fixture(scope="session")
def oracle_big_bunny(built, big_bunny_inputs):
    from oracle import oracle as O
    v, c, _ = big_bunny_inputs
    return O.OracleScene(v, c, True)
