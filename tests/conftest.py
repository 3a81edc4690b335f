import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason='no CUDA device')
    for item in items:
        if 'gpu' in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope='session')
def golden_dir():
    return GOLDEN


@pytest.hookimpl(hookwrapper=True)
def pytest_runtest_makereport(item, call):
    """A GPU test that dies on a CUDA launch failure: attach where the kernel's bounded wait gave up
    (include/nsg_b200.h: nsg_hang_report) to the report."""
    outcome = yield
    rep = outcome.get_result()
    if rep.when == 'call' and rep.failed and 'gpu' in item.keywords:
        try:
            from neural_sound_generation_b200 import _cabi
            r = _cabi.hang_report()
            if r:
                rep.sections.append(('nsg hang report', r))
                print('\n[nsg hang report] ' + r)
        except Exception:
            pass
