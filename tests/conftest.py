import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


import clientsim as cs  # noqa: E402  (test infrastructure: synthetic herd-client keys and ciphertexts)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import pyoracle
    pyoracle.lib()
    return pyoracle


@pytest.fixture(scope="session")
def test_key():
    """n = 32 test parameter set: same ring / moduli / decomposition as STD128, 16x fewer CMUX steps"""
    from herdsman_b200 import api
    sk, skr, blob = cs.keygen(api.N_TEST, seed=0x48455244)
    return {"n": api.N_TEST, "sk": sk, "skr": skr, "blob": blob}


@pytest.fixture(scope="session")
def std128_key():
    from herdsman_b200 import api
    sk, skr, blob = cs.keygen(api.N_STD128, seed=0x48455244)
    return {"n": api.N_STD128, "sk": sk, "skr": skr, "blob": blob}
