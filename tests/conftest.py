"""pytest configuration: registers the `gpu` marker, imports the hyphenated package dir, and
builds the oracle (gcc) + the CUDA library (nvcc cross-compiles without a GPU) once per session."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import __graft_entry__ as ge  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    return ge.load_package()


@pytest.fixture(scope="session")
def orc():
    ge.build_oracle()
    from oracle import oracle
    oracle.lib()
    return oracle


@pytest.fixture(scope="session")
def built_lib():
    ge.build_library()
    return ge.LIB


def _has_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def engine(pkg, built_lib):
    if not _has_gpu():
        pytest.skip("no CUDA device")
    with pkg.Engine(device=0) as e:
        yield e


@pytest.fixture
def late_reorder(engine, pkg):
    """igs_append_events accepts late events (ordered by timestamp) for the duration of one test."""
    engine.set_late_policy(pkg.abi.LATE_REORDER)
    yield engine
    engine.set_late_policy(pkg.abi.LATE_REJECT)
