"""`PixelRayDataset`: the reference's per-pixel dataset (nerf/dataset.py:175-316) with the same
item dictionary, plus a batched GPU gatherer (`gather`, `batches`) that replaces 1024 Python
`__getitem__` calls + default collate per batch by ONE kernel launch.

Constructor generations:
  G1  PixelRayDataset(images, poses[N,4,4], focal_length)                     (example.ipynb cell 5)
  G2  PixelRayDataset(images, poses[N,3,4], focal_length, states_x=...)       (train_conditional_nerf.py:97)
  HEAD PixelRayDataset(images, segmentation, poses, focal_length, ...)        (nerf/dataset.py:194)
The HEAD form is recognised by its 4th positional being the focal length (a number).
"""
import numbers

import torch
import torch.utils.data as data

from . import ops


class PixelRayDataset(data.Dataset):

    def __init__(self, images, *args, states_x=None, states_d=None, x_states=None, d_states=None,
                 segmentation=None, poses=None, focal_length=None):
        # leading tensors before the focal length: (poses) = G1/G2, (segmentation, poses) = HEAD; each of
        # segmentation / poses / focal_length may also come as a keyword (nerf/dataset.py:194-195)
        rest, lead = list(args), []
        while rest and len(lead) < 2 and not isinstance(rest[0], numbers.Number):
            lead.append(rest.pop(0))
        if rest and isinstance(rest[0], numbers.Number) and focal_length is None:
            focal_length = rest.pop(0)
        if poses is None and lead:
            poses = lead.pop()
        if segmentation is None and lead:
            segmentation = lead.pop()
        if poses is None or focal_length is None or lead:
            raise TypeError("PixelRayDataset(images, poses, focal_length, ...) or PixelRayDataset(images, segmentation, poses, focal_length, ...)")
        if len(rest) > 0 and states_x is None:
            states_x = rest[0]
        if len(rest) > 1 and states_d is None:
            states_d = rest[1]
        self.images = images
        self.segmentation = segmentation
        self.poses = poses
        self.focal_length = float(focal_length)
        self.states_x = states_x if states_x is not None else x_states
        self.states_d = states_d if states_d is not None else d_states
        self.x_states, self.d_states = self.states_x, self.states_d     # attribute names seen in the notebook traceback
        self._rays = None

    @property
    def rays(self):
        """Camera-space rays [H,W,3] (reference: built once in __init__, nerf/dataset.py:228)."""
        if self._rays is None:
            h, w = self.images.shape[1], self.images.shape[2]
            if self.images.is_cuda:
                from .model import NeRF
                self._rays = NeRF.generate_rays(h, w, self.focal_length, dtype=self.images.dtype, device=self.images.device)
            else:       # host tensors: same arithmetic in torch (3 elementwise ops on an [H,W] grid)
                y, x = torch.meshgrid(torch.arange(h, dtype=self.images.dtype), torch.arange(w, dtype=self.images.dtype), indexing="ij")
                x = (x - 0.5 * float(w - 1)) / self.focal_length
                y = (y - 0.5 * float(h - 1)) / self.focal_length
                self._rays = torch.stack([x, -y, -torch.ones_like(x)], dim=-1)
        return self._rays

    def __len__(self):
        return self.images.shape[0] * self.images.shape[1] * self.images.shape[2]

    @staticmethod
    def decode_index(idx, n, h, w):
        """idx -> (image_bi, image_hi, image_wi), reference nerf/dataset.py:283-291."""
        wi = idx % w; idx = idx // w
        hi = idx % h; idx = idx // h
        return idx % n, hi, wi

    def __getitem__(self, idx):
        """Per-item path kept for API parity (DataLoader users); same keys as the reference,
        including its `pose_d = rays_d` quirk (nerf/dataset.py:315)."""
        n, h, w = self.images.shape[:3]
        bi, hi, wi = self.decode_index(int(idx), n, h, w)
        pixel = self.images[bi, hi, wi]
        ray = self.rays[hi, wi]
        pose = self.poses[bi]
        dev = pixel.device
        state_x = self.states_x[bi] if self.states_x is not None else torch.zeros(0, device=dev)
        state_d = self.states_d[bi] if self.states_d is not None else torch.zeros(0, device=dev)
        pose_o, pose_r = pose[:3, 3], pose[:3, :3]
        rays_d = (pose_r * ray.unsqueeze(-2)).sum(dim=-1)
        item = dict(image_wi=torch.LongTensor([wi]).to(dev), image_hi=torch.LongTensor([hi]).to(dev),
                    image_bi=torch.LongTensor([bi]).to(dev), states_x=state_x, states_d=state_d,
                    pixels=pixel, rays=ray, pose_o=pose_o, pose_d=rays_d, rays_o=pose_o, rays_d=rays_d)
        if self.segmentation is not None:
            item["label"] = self.segmentation[bi, hi, wi]
        return item

    # ---- batched path: one kernel per batch ---------------------------------------------
    def gather(self, idx):
        """Whole batch at once on the GPU: same dictionary as default-collating __getitem__."""
        idx = idx.to(self.images.device)
        o, d, px, st, bhw = ops.gather_pixel_rays(self.images, self.poses, self.focal_length, idx, self.states_x)
        b = idx.numel()
        out = dict(image_bi=bhw[:, 0:1], image_hi=bhw[:, 1:2], image_wi=bhw[:, 2:3], pixels=px, rays_o=o, rays_d=d,
                   pose_o=o, pose_d=d, states_x=st if st is not None else torch.zeros(b, 0, device=o.device),
                   states_d=(self.states_d[bhw[:, 0]] if self.states_d is not None else torch.zeros(b, 0, device=o.device)))
        out["rays"] = self.rays[bhw[:, 1], bhw[:, 2]]
        if self.segmentation is not None:
            out["label"] = self.segmentation[bhw[:, 0], bhw[:, 1], bhw[:, 2]]
        return out

    def batches(self, batch_size, shuffle=True, generator=None, rank=0, world_size=1, drop_last=False):
        """DataLoader(dataset, batch_size, shuffle) replacement: yields gathered batches.  With
        world_size > 1 every rank draws the same permutation and takes its strided share."""
        n = len(self)
        dev = self.images.device
        order = torch.randperm(n, generator=generator) if shuffle else torch.arange(n)
        # every rank must run the same number of steps with the same batch sizes (each step is one all-reduce):
        # cut the permutation to a multiple of world_size before striding, like DistributedSampler(drop_last=True)
        order = order[:(n // world_size) * world_size]
        order = order[rank::world_size].to(dev)
        for i in range(0, order.numel(), batch_size):
            sel = order[i:i + batch_size]
            if drop_last and sel.numel() < batch_size:
                break
            yield self.gather(sel)


class ImageRayDataset(data.Dataset):
    """The reference's block-stratified dataset (nerf/dataset.py:6-172): every item draws
    `num_samples_per_block` pixels from each cell of a `num_vertical_blocks x num_horizontal_blocks`
    grid over ONE image and returns them as [num_samples_per_block, blocks, ...] tensors with the
    reference's keys.  The index arithmetic (:134-153) is a handful of integer ops on a
    [samples, blocks] grid and stays in torch; pixels, poses and world-space rays of the whole item come
    from ONE launch of the batched gather kernel (`nerf_gather_pixel_rays`) when the images live on the
    GPU.  `block_ray_idx` may be passed in (tests, reproducible runs); by default it is drawn like the
    reference does (`torch.multinomial` over uniform weights, with replacement)."""

    def __init__(self, images, poses, states, focal_length, num_vertical_blocks=8, num_horizontal_blocks=8,
                 num_samples_per_block=2):
        self.images, self.poses, self.states = images, poses, states
        self.focal_length = float(focal_length)
        self.num_vertical_blocks, self.num_horizontal_blocks = num_vertical_blocks, num_horizontal_blocks
        self.num_samples_per_block = num_samples_per_block
        self.horizontal_block_size = images.shape[2] // num_horizontal_blocks      # :70-73
        self.vertical_block_size = images.shape[1] // num_vertical_blocks
        self.total_block_area = self.horizontal_block_size * self.vertical_block_size
        self._rays = None

    rays = PixelRayDataset.rays

    def __len__(self):
        return (self.images.shape[0] * self.total_block_area) // self.num_samples_per_block     # :93-94

    def block_pixels(self, idx, block_ray_idx=None, generator=None):
        """(image_bi, image_hi, image_wi), each [samples, blocks] int64 -- nerf/dataset.py:134-158."""
        dev = self.images.device
        blocks = self.num_vertical_blocks * self.num_horizontal_blocks
        if block_ray_idx is None:
            block_ray_idx = torch.multinomial(torch.ones(self.num_samples_per_block, self.total_block_area), blocks,
                                              replacement=True, generator=generator)
        block_ray_idx = block_ray_idx.to(dev)
        block_idx = torch.arange(blocks, device=dev).unsqueeze(0)
        image_wi = block_ray_idx % self.horizontal_block_size + (block_idx % self.num_horizontal_blocks) * self.horizontal_block_size
        image_hi = block_ray_idx // self.horizontal_block_size + (block_idx // self.num_horizontal_blocks) * self.vertical_block_size
        image_bi = torch.full_like(image_hi, int(idx) // self.total_block_area)
        return image_bi, image_hi, image_wi

    def __getitem__(self, idx, block_ray_idx=None, generator=None):
        bi, hi, wi = self.block_pixels(idx, block_ray_idx, generator)
        shape = hi.shape
        h, w = self.images.shape[1], self.images.shape[2]
        pose = self.poses[bi]
        pose_o, pose_d = pose[..., :3, 3], pose[..., :3, :3]
        if self.images.is_cuda:
            flat = ((bi * h + hi) * w + wi).reshape(-1)
            o, d, px, _, _ = ops.gather_pixel_rays(self.images, self.poses, self.focal_length, flat)
            pixel, rays_o, rays_d = px.view(*shape, -1), o.view(*shape, 3), d.view(*shape, 3)
            ray = self.rays[hi, wi]
        else:                                                       # host tensors: the reference's own indexing
            pixel = self.images[bi, hi, wi]
            ray = self.rays[hi, wi]
            rays_o, rays_d = pose_o, (pose_d * ray.unsqueeze(-2)).sum(dim=-1)
        return dict(image_bi=bi, image_hi=hi, image_wi=wi, pixels=pixel, states=self.states[bi], rays=ray,
                    pose_o=pose_o, pose_d=pose_d, rays_o=rays_o, rays_d=rays_d)
