import os
import sys
import warnings

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
warnings.filterwarnings("ignore", category=FutureWarning)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def spec():
    from nerf_emitter_b200.params import EmitterModelSpec
    return EmitterModelSpec()


_PARAM_CACHE = {}


@pytest.fixture(scope="session")
def params_factory(spec):
    from nerf_emitter_b200.params import make_params

    def get(seed=0, table_scale=1e3):
        key = (seed, table_scale)
        if key not in _PARAM_CACHE:
            _PARAM_CACHE[key] = make_params(spec, seed=seed, table_scale=table_scale)
        return _PARAM_CACHE[key]
    return get
