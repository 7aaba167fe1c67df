import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session')
def orc():
    """The CPU oracle (test infrastructure, oracle/orc.py)."""
    from oracle import orc as _orc
    _orc.lib()
    return _orc


def to_oracle_scene(orc_mod, scene):
    """Product Scene -> oracle Scene (both mirror include/mg_scene.h byte for byte)."""
    import ctypes as C
    out = orc_mod.Scene()
    assert C.sizeof(out) == C.sizeof(scene)
    C.memmove(C.byref(out), C.byref(scene), C.sizeof(scene))
    return out
