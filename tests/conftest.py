import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


from canopy_b200 import _lib as _canopy_lib  # noqa: E402

_canopy_lib.use_testing_library()   # the suite drives the test hooks of include/canopy-testing.h; the product library has none


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build libcanopy.so and the oracle once per session (no-op when up to date)."""
    import __graft_entry__ as ge
    ge.build()


def pytest_generate_tests(metafunc):
    """Every GPU test runs twice: with the engine chosen by the library (the fused shared-memory kernel for models of up to
    about 200 live nodes, which most test graphs are) and with the streamed sampler + gate-walk pair forced, so that
    neither path hides behind the other."""
    if metafunc.definition.get_closest_marker("gpu") is not None:
        metafunc.fixturenames.append("_engine_mode")
        metafunc.parametrize("_engine_mode", ["auto", "streamed"], indirect=True)


@pytest.fixture
def _engine_mode(request):
    from canopy_b200 import engine
    old = engine.DEFAULT_ENGINE
    engine.DEFAULT_ENGINE = {"auto": engine.ENGINE_AUTO, "streamed": engine.ENGINE_STREAMED}[request.param]
    yield request.param
    engine.DEFAULT_ENGINE = old
