"""nerf_b200: B200-native NeRF hot path (ray generation, sampling, encoding, coordinate MLP,
volume rendering; forward and backward) behind the API of brandontrabucco/nerf."""
from .model import NeRF  # noqa: F401
from .dataset import PixelRayDataset  # noqa: F401
