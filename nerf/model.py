"""`from nerf.model import NeRF` -- the B200-native module behind the reference's import path."""
from nerf_b200.model import NeRF  # noqa: F401
