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


@pytest.fixture(scope="session")
def tiny_galaxy():
    """Galaxy-shaped grid (C3), max_lvl 2: 757 cells, 231 emitting leaf cells, 4 wavelengths, z wall."""
    from dartray_b200 import scenes
    return scenes.galaxy_scene(max_lvl=2, min_lvl=2, lnum=4, name="galaxy_tiny")


@pytest.fixture(scope="session")
def small_galaxy():
    """Galaxy-shaped grid (C3), 3 levels: 3322 cells, 2909 emitting leaf cells, 3 wavelengths."""
    from dartray_b200 import scenes
    return scenes.galaxy_scene(max_lvl=3, min_lvl=1, lnum=3, max_dtau=0.3, max_dlum=2e-3, name="galaxy_small")


@pytest.fixture(scope="session")
def small_magtar():
    """MAGTAR ellipsoidal shell (C2), max_lvl 2: 2513 cells, point source at the origin, kp_sca in {-1,0,1}."""
    from dartray_b200 import scenes
    return scenes.magtar_scene(max_lvl=2, lnum=4, name="magtar_small")
