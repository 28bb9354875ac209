import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box)")


@pytest.fixture
def vdp_cfg():
    """Set process-wide switches of the CUDA library for one test (``vdp_cfg(SWEEPS="tiled", NSEG=3)``, values as the
    VDP_<KEY> environment variables take them); every switch is back at its default when the test ends."""
    from vi_diffusion_processes_b200 import _lib

    def setter(**kw):
        conv = {}
        for k, v in kw.items():
            if v is not None and k.upper() != "SWEEPS":
                v = int(v)
            conv[k] = v
        _lib.configure(**conv)
    yield setter
    _lib.configure(**{k: None for k in _lib.CONFIG_KEYS})
