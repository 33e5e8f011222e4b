"""Drop-in import path of the reference package (`from nerf.model import NeRF`,
`from nerf.dataset import PixelRayDataset`): thin re-exports of nerf_b200."""
