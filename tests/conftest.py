import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def small_trusti():
    """TRUST-I slab, 4 wavelengths, max_lvl 3: 5801 cells, 1 point source, 7 directions."""
    from dartray_b200 import scenes
    return scenes.trusti_scene(lnum=4, max_lvl=3)


@pytest.fixture(scope="session")
def tiny_trusti():
    """TRUST-I slab, 3 wavelengths, max_lvl 2 (1801 cells), optically thick (tau_z = 1)."""
    from dartray_b200 import scenes
    return scenes.trusti_scene(tau_z=1.0, base=(10, 2), max_lvl=2, lnum=3, name="trustI_tiny")
