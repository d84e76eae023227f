import os
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)
GOLDEN = os.path.join(REPO, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def lib():
    import peasoup_b200
    if not os.path.exists(peasoup_b200.LIB_PATH):
        peasoup_b200.build()
    return peasoup_b200.load()


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.lib()
    return O


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference kernels (oracle/_ref/libpeasoup_ref.so), GPU box only."""
    import ctypes
    path = os.path.join(REPO, "oracle", "_ref", "libpeasoup_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libpeasoup_ref.so not built (make -C oracle ref)")
    r = ctypes.CDLL(path)
    r.ref_last_error.restype = ctypes.c_char_p
    r.ref_worker_search.restype = ctypes.c_long
    return r


@pytest.fixture(scope="session")
def tutorial_trials(oracle):
    """tutorial.fil dedispersed to the golden run's 59 DM trials by the oracle's dedisp
    restatement (cached in /tmp: ~20 s of numpy)."""
    ov = oracle.parse_overview_xml(os.path.join(GOLDEN, "overview.xml"))
    cache = "/tmp/peasoup_b200_tutorial_trials.npy"
    if os.path.exists(cache):
        trials = np.load(cache)
    else:
        hdr, data = oracle.read_sigproc(os.path.join(GOLDEN, "tutorial.fil"))
        trials, _ = oracle.dedisperse(data, hdr, ov["dm_list"])
        np.save(cache, trials)
    return trials, ov
