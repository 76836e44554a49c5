import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)


def pytest_configure(config):
  config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


def _make(directory, target):
  subprocess.check_call(['make', '-C', directory, target], stdout=subprocess.DEVNULL)


@pytest.fixture(scope='session', autouse=True)
def _build_test_libs():
  """Build the checker libraries (oracle, host build of the core).  The product
  library is built by __graft_entry__.build(), never here."""
  _make(os.path.join(ROOT, 'oracle'), 'libsolo_oracle.so')
  _make(os.path.join(ROOT, 'tests', 'hostsim'), 'libhostsim.so')
