"""Shared fixtures.  `-m gpu` tests need a B200 and call the product through the C ABI (libggq.so);
everything else runs on CPU.  oracle/ is imported here and ONLY here / in tests (it is the checker)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def _has_gpu():
    try:
        import genome_graph_annotation_b200 as g
        return g.device_count() > 0
    except Exception:
        return False


HAS_GPU = None


def pytest_collection_modifyitems(config, items):
    global HAS_GPU
    if HAS_GPU is None:
        HAS_GPU = _has_gpu()
    if HAS_GPU:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference compiled into oracle/_ref/libggref.so."""
    from oracle import ref_binding
    if not ref_binding.available():
        pytest.skip("oracle/_ref/libggref.so not built (needs /root/reference; run `make -C oracle`)")
    return ref_binding


@pytest.fixture(scope="session")
def orc():
    """The plain-C restatement oracle/liboracle_c.so."""
    from oracle import c_binding
    if not c_binding.available():
        pytest.skip("oracle/liboracle_c.so not built (run `make -C oracle restatement`)")
    return c_binding


@pytest.fixture(scope="session")
def ggq():
    import genome_graph_annotation_b200 as g
    g.lib()
    return g


def random_dense(n, m, density, seed):
    rng = np.random.default_rng(seed)
    return (rng.random((n, m)) < density).astype(np.uint8)
