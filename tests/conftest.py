import os
import sys
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (HERE, ROOT, os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


# The driver runs `pytest -m gpu -x`: the tests that compare the CUDA path with the oracle / the
# reference's golden outputs must run first, the multi-GB full-size cases last, so that a failure
# late in the run never hides the parity evidence.
_ORDER = ["test_oracle", "test_host_cpu", "test_ms_fast_host", "test_ms_fast_repeats", "test_multi_gloo",
          "test_gpu_parity", "test_gpu_ordering", "test_gpu_escapes", "test_gpu_sharded", "test_gpu_consumers",
          "test_gpu_fullsize", "test_zz_no_alloc"]


def pytest_collection_modifyitems(session, config, items):
    def key(it):
        name = os.path.splitext(os.path.basename(str(it.fspath)))[0]
        return _ORDER.index(name) if name in _ORDER else len(_ORDER) - 2      # test_zz_no_alloc stays last
    items.sort(key=key)                # stable: the order inside a file is kept


@pytest.fixture(scope="session")
def workdir(tmp_path_factory):
    return str(tmp_path_factory.mktemp("cmsbwt"))
