"""Golden items of the UNMODIFIED reference `nerf.dataset.ImageRayDataset` (run in the build container,
where /root/reference exists):  python tests/golden/make_image_ray_dataset.py
Inputs are regenerated from the seeds below by the test; the item's random block offsets come from the
global torch generator, seeded right before `__getitem__`."""
import os, sys, warnings
import numpy as np, torch
warnings.filterwarnings("ignore")
sys.path.insert(0, "/root/reference")
from nerf.dataset import ImageRayDataset  # noqa: E402


def inputs():
    g = torch.Generator().manual_seed(11)
    images = torch.rand(3, 16, 24, 3, generator=g)
    poses = torch.eye(4).repeat(3, 1, 1); poses[:, :3, :] = torch.randn(3, 3, 4, generator=g)
    states = torch.randn(3, 5, generator=g)
    return images, poses, states


if __name__ == "__main__":
    images, poses, states = inputs()
    ds = ImageRayDataset(images, poses, states, 20.0, num_vertical_blocks=4, num_horizontal_blocks=4, num_samples_per_block=2)
    out = {"len": np.int64(len(ds))}
    for idx in (0, 17, 35):
        torch.manual_seed(100 + idx)
        item = ds[idx]
        for k, v in item.items():
            out[f"{idx}/{k}"] = v.numpy()
    np.savez_compressed(os.path.join(os.path.dirname(__file__), "image_ray_dataset.npz"), **out)
    print({k: v.shape for k, v in out.items() if k.startswith("17/")})
