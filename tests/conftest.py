import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_table():
    return np.load(os.path.join(GOLDEN, "dist_table.npz"))["sm"]


@pytest.fixture(scope="session")
def table():
    return load_table()


@pytest.fixture(scope="session")
def oracle(table):
    from oracle.pyoracle import Oracle
    return Oracle(table)


@pytest.fixture(scope="session")
def reference():
    """The unmodified reference functions; only where oracle/_ref was built."""
    from oracle.pyoracle import Reference, have_reference
    if not have_reference():
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    return Reference(_dst_file())


_DST = None


def _dst_file():
    """eucledian.dst rebuilt from the golden fixture (header 777,10,5 + table)."""
    global _DST
    if _DST is None:
        import tempfile
        d = tempfile.mkdtemp(prefix="vsk_dst_")
        _DST = os.path.join(d, "eucledian.dst")
        with open(_DST, "wb") as f:
            f.write(np.array([777, 10, 5], np.int32).tobytes())
            f.write(load_table().tobytes())
    return _DST


@pytest.fixture(scope="session")
def dst_file():
    return _dst_file()


@pytest.fixture(scope="session")
def ctx(table):
    """A libvsk context on cuda:0 (gpu tests only). Fails loudly without the library."""
    import vntrseek_b200 as v
    c = v.Context(table, 0)
    yield c
    c.close()


def golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=True)
