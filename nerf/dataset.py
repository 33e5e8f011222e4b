"""`from nerf.dataset import PixelRayDataset, ImageRayDataset` -- the reference's import path."""
from nerf_b200.dataset import ImageRayDataset, PixelRayDataset  # noqa: F401
