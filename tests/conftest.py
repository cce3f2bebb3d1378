import os
import subprocess
import sys

import numpy
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


def golden(name):
    return numpy.load(os.path.join(GOLDEN, name + '.npz'))


@pytest.fixture(scope='session')
def pins():
    import json
    with open(os.path.join(GOLDEN, 'pins.json')) as fh:
        return json.load(fh)


@pytest.fixture(scope='session')
def oracle_lib():
    """The CPU oracle behind the product's C ABI (test infrastructure)."""
    from atooms_b200 import capi
    path = os.path.join(ROOT, 'oracle', '_build', 'liboracle_b2.so')
    src = os.path.join(ROOT, 'oracle', 'oracle_c.c')
    if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
        subprocess.check_call(['make', '-C', os.path.join(ROOT, 'oracle')])
    return capi.Library(path)


@pytest.fixture(scope='session')
def cuda_lib():
    """The product library; GPU tests fail loudly if it is missing."""
    from atooms_b200 import capi
    return capi.load()


def make_engine(lib, data, table, precision=64, skin=0.3, velocities=True):
    from atooms_b200 import capi
    eng = capi.Engine(precision, library=lib)
    eng.set_box(data['side'], data['center'])
    eng.set_pair_table(table.eps, table.sigma, table.rcut, table.shifted)
    eng.set_skin(skin)
    eng.upload(data['position'], data['species'],
               data['velocity'] if velocities and 'velocity' in data else None,
               data['mass'] if 'mass' in data else None)
    return eng
