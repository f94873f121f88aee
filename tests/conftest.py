import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_traces():
    import gzip
    import json

    with gzip.open(os.path.join(ROOT, "tests", "golden", "traces.json.gz"), "rt") as f:
        return json.load(f)


def _backend_params():
    return [
        pytest.param("oracle", id="oracle"),
        pytest.param("hostsim", id="hostsim"),
        pytest.param("cuda", id="cuda", marks=pytest.mark.gpu),
    ]


@pytest.fixture(params=_backend_params())
def backend(request):
    """Engine under test: the oracle and the host build of the device headers run on CPU; "cuda" is
    the product (libmrts_b200.so through MicroRTSGridModeVecEnv) and needs the B200."""
    return request.param


@pytest.fixture(scope="session")
def known_answers():
    import json

    with open(os.path.join(ROOT, "tests", "golden", "known_answers.json")) as f:
        return json.load(f)
