import json
import os
import sys
import tarfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (B200); run with -m gpu')


@pytest.fixture(scope='session')
def checker():
    """The CPU checkers (oracle/): built on demand; test infrastructure only."""
    from oracle import checker as ck
    ck.build()
    return ck


@pytest.fixture(scope='session')
def ref_fastqs(tmp_path_factory):
    """The reference's own FastQ fixtures, unpacked from tests/golden/ref_inputs.tar.gz."""
    d = tmp_path_factory.mktemp('ref_fastqs')
    with tarfile.open(os.path.join(GOLDEN, 'ref_inputs.tar.gz')) as tar:
        tar.extractall(str(d), filter='data')
    return str(d)


@pytest.fixture(scope='session')
def golden():
    return json.load(open(os.path.join(GOLDEN, 'golden.json')))


@pytest.fixture(scope='session')
def gpu_engine():
    """kvarq_b200.engine on cuda:0 -- raises (never falls back) when no GPU is there."""
    from kvarq_b200 import engine
    engine.get_config()
    return engine
