import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


def golden_names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, '*.npz')))


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + '.npz'), allow_pickle=False)
    return {k: z[k] for k in z.files}


# goldens whose kernel has a closed form (the device path covers these)
CLOSED_FORM = [n for n in golden_names() if '_m19_' not in n]


@pytest.fixture(scope='session')
def lib():
    """The C-ABI library; GPU tests call the product through it."""
    from polyaxon_b200 import _lib
    return _lib.load()
