"""The oracle's known-answer tests and the host-builder / flatten checks once more under the `gpu` marker, so that the
driver's `pytest -m gpu` run on the B200 box also shows the oracle pinned THERE (same g++, same build of the checker)
-- the parity tests of test_gpu_parity.py are only as good as the oracle they compare with."""
import pytest

from tests import test_host_builder as _H
from tests import test_oracle_kat as _K

pytestmark = pytest.mark.gpu

for _mod in (_K, _H):
    for _name in dir(_mod):
        if _name.startswith("test_"):
            globals()["test_box_" + _name[5:]] = getattr(_mod, _name)
del _mod, _name
