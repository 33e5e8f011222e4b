"""`from nerf.dataset import PixelRayDataset` -- the reference's import path."""
from nerf_b200.dataset import PixelRayDataset  # noqa: F401
