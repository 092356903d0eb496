import os
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle_mod():
    from oracle import oracle as O
    O.lib()
    return O


_KEYS = {}


@pytest.fixture(scope="session")
def keys_factory(oracle_mod):
    def get(name, seed=3):
        if (name, seed) not in _KEYS:
            _KEYS[(name, seed)] = oracle_mod.Keys(name, seed=seed)
        return _KEYS[(name, seed)]
    return get


_ENGINES = {}


@pytest.fixture(scope="session")
def engine_factory(keys_factory):
    """GPU engine with the oracle-generated (seeded) server keys loaded."""
    def get(name, seed=3):
        if (name, seed) not in _ENGINES:
            import helm_b200
            K = keys_factory(name, seed)
            eng = helm_b200.Engine(K.params, device=0)
            eng.load_keys(K.bsk(), K.ksk())
            _ENGINES[(name, seed)] = eng
        return _ENGINES[(name, seed)]
    return get
