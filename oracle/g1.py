"""G1 ("nerf-pytorch v1.2") oracle -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Plain eager-PyTorch fp32 restatement of the NeRF generation that
``/root/reference/examples/nerf.pth`` was saved from and that
``/root/reference/examples/example.ipynb`` (cells 5-9) and
``/root/reference/train_conditional_nerf.py:86-150`` call.  Its source is not in
the reference tree; the restatement follows SURVEY.md Appendix A.3 and re-uses,
line for line in meaning, the pieces of HEAD ``nerf/model.py`` that did not
change between generations:

* ray generation            -> /root/reference/nerf/model.py:242-278
* camera->world transform   -> /root/reference/nerf/model.py:336-367
* stratified jitter         -> /root/reference/nerf/model.py:422-433
* positional-encoding layout-> /root/reference/nerf/model.py:216-240 (dead code, layout only)
* compositing weights       -> /root/reference/nerf/model.py:437-469
* colour composite          -> /root/reference/nerf/model.py:660
* chunked image render      -> /root/reference/nerf/model.py:727-770

PARITY: pinned by the checkpoint structure + the notebook's embedded render
(tests/test_oracle.py); bit-level details are "parity unpinned" (SURVEY A.4).
"""
import math

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as functional

PE_X_OCTAVES = 10   # block_0.0.weight is [256, 60] = 3 coords * 2 * 10
PE_D_OCTAVES = 6    # block_2.0.weight is [256, 256 + 36], 36 = 3 * 2 * 6
HIDDEN = 256


def positional_encoding(v, octaves):
    """sin/cos features, layout of /root/reference/nerf/model.py:235-240:
    per coordinate ``[sin(v f_0..f_{L-1}) | cos(v f_0..f_{L-1})]``, then flatten.
    Frequencies f_j = pi * 2^j."""
    # exact powers of two (torch.pow is not exact on every CPU), times float32(pi)
    freq = torch.tensor([2.0 ** j for j in range(octaves)], dtype=v.dtype, device=v.device) * np.pi
    arg = v.unsqueeze(-1) * freq
    emb = torch.cat([torch.sin(arg), torch.cos(arg)], dim=-1)
    return torch.flatten(emb, start_dim=-2, end_dim=-1)


def generate_rays(image_h, image_w, focal_length, dtype=torch.float32, device="cpu"):
    """/root/reference/nerf/model.py:242-278 (meshgrid 'ij': y=row, x=col)."""
    kwargs = dict(dtype=dtype, device=device)
    y, x = torch.meshgrid(torch.arange(image_h, **kwargs),
                          torch.arange(image_w, **kwargs), indexing="ij")
    x = (x - 0.5 * float(image_w - 1)) / focal_length
    y = (y - 0.5 * float(image_h - 1)) / focal_length
    return torch.stack([x, -y, -torch.ones_like(x)], dim=-1)


def rays_to_world_coordinates(rays, camera_o, camera_r):
    """/root/reference/nerf/model.py:367."""
    return camera_o, (camera_r * rays.unsqueeze(-2)).sum(dim=-1)


def stratified_t(near, far, num_samples, batch_shape, u=None,
                 dtype=torch.float32, device="cpu"):
    """G1 sampler: linspace(near, far, S) + the jitter of
    /root/reference/nerf/model.py:422-433.  ``u`` (uniforms in [0,1), shape
    batch_shape+[S]) is passed in so the CUDA path can be fed the same draws."""
    t = torch.linspace(near, far, num_samples, dtype=dtype, device=device)
    t = torch.broadcast_to(t.reshape([1] * len(batch_shape) + [num_samples]),
                           list(batch_shape) + [num_samples])
    if u is not None:
        mid = 0.5 * (t[..., 1:] + t[..., :-1])
        lower = torch.cat([t[..., :1], mid], dim=-1)
        upper = torch.cat([mid, t[..., -1:]], dim=-1)
        t = lower + (upper - lower) * u
    return t


def alpha_compositing_coefficients(points, density_outputs):
    """/root/reference/nerf/model.py:462-469, verbatim in meaning."""
    dists = points[..., 1:, :] - points[..., :-1, :]
    dists = functional.pad(torch.linalg.norm(dists, dim=-1, keepdim=True),
                           (0, 0, 0, 1), value=1e10)
    alpha = torch.exp(-functional.relu(density_outputs) * dists)
    return (1.0 - alpha) * functional.pad(
        torch.cumprod(alpha[..., :-1, :] + 1e-10, dim=-2), (0, 0, 1, 0), value=1.0)


def sample_pdf(t, weights, num_fine, u=None):
    """Hierarchical (inverse-CDF) sampler of the NeRF paper, section 5.2.
    NOT in any reference source (only promised at model.py:191-193) -- this is
    our own restatement, "parity unpinned".  Bins are the midpoints of the
    coarse ``t``; weights of the interior samples (+1e-5) form the pdf.
    ``u`` defaults to the deterministic linspace(0,1,num_fine)."""
    mids = 0.5 * (t[..., 1:] + t[..., :-1])                     # [B, S-1]
    w = weights[..., 1:-1] + 1e-5                               # [B, S-2]
    pdf = w / w.sum(dim=-1, keepdim=True)
    cdf = torch.cumsum(pdf, dim=-1)
    cdf = torch.cat([torch.zeros_like(cdf[..., :1]), cdf], dim=-1)   # [B, S-1]
    if u is None:
        u = torch.linspace(0.0, 1.0, num_fine, dtype=t.dtype, device=t.device)
        u = torch.broadcast_to(u, list(cdf.shape[:-1]) + [num_fine])
    u = u.contiguous()
    idx = torch.searchsorted(cdf.contiguous(), u, right=True)
    below = torch.clamp(idx - 1, min=0)
    above = torch.clamp(idx, max=cdf.shape[-1] - 1)
    cdf_b = torch.gather(cdf, -1, below)
    cdf_a = torch.gather(cdf, -1, above)
    bin_b = torch.gather(mids, -1, below)
    bin_a = torch.gather(mids, -1, above)
    denom = cdf_a - cdf_b
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    return bin_b + (u - cdf_b) / denom * (bin_a - bin_b)


def _unit(n_in, n_out):
    return [nn.Linear(n_in, n_out), nn.ReLU(), nn.LayerNorm(n_out)]


class G1NeRF(nn.Module):
    """state_dict key set == examples/nerf.pth: heads first, then
    block_0/1/2 as Sequential(Linear, ReLU, LayerNorm) units."""

    def __init__(self, normalize_position=1.0, density_inputs=3, color_inputs=3,
                 hidden_size=HIDDEN, normalize_direction=True):
        super().__init__()
        self.normalize_position = float(normalize_position)
        self.density_inputs = density_inputs
        self.color_inputs = color_inputs
        self.hidden_size = hidden_size
        self.normalize_direction = normalize_direction   # A.4 switch (1)
        ex = 2 * PE_X_OCTAVES * density_inputs
        ed = 2 * PE_D_OCTAVES * color_inputs
        h = hidden_size
        self.density = nn.Linear(h, 1)
        self.color = nn.Linear(h, 3)
        self.block_0 = nn.Sequential(*(_unit(ex, h) + _unit(h, h) + _unit(h, h) + _unit(h, h)))
        self.block_1 = nn.Sequential(*(_unit(h + ex, h) + _unit(h, h) + _unit(h, h) + _unit(h, h)))
        self.block_2 = nn.Sequential(*(_unit(h + ed, h) + _unit(h, h)))

    # -- per-point network -------------------------------------------------
    def encode(self, rays_o, rays_d, t, states_x=None, states_d=None):
        pts = rays_o.unsqueeze(-2) + rays_d.unsqueeze(-2) * t.unsqueeze(-1)
        x = pts / self.normalize_position
        if states_x is not None and states_x.shape[-1] > 0:
            x = torch.cat([x, torch.broadcast_to(
                states_x.unsqueeze(-2), list(x.shape[:-1]) + [states_x.shape[-1]])], dim=-1)
        d = rays_d
        if self.normalize_direction:
            d = d / torch.linalg.norm(d, dim=-1, keepdim=True)
        if states_d is not None and states_d.shape[-1] > 0:
            d = torch.cat([d, states_d], dim=-1)
        ex = positional_encoding(x, PE_X_OCTAVES)
        ed = positional_encoding(d, PE_D_OCTAVES)
        return pts, ex, ed

    def mlp(self, ex, ed, return_activations=False):
        acts = []
        h = ex
        for i in range(0, 12, 3):
            h = self.block_0[i + 2](self.block_0[i + 1](self.block_0[i](h)))
            acts.append(h)
        h = torch.cat([h, ex], dim=-1)
        for i in range(0, 12, 3):
            h = self.block_1[i + 2](self.block_1[i + 1](self.block_1[i](h)))
            acts.append(h)
        density = self.density(h)
        edb = torch.broadcast_to(ed.unsqueeze(-2), list(h.shape[:-1]) + [ed.shape[-1]])
        g = torch.cat([h, edb], dim=-1)
        for i in range(0, 6, 3):
            g = self.block_2[i + 2](self.block_2[i + 1](self.block_2[i](g)))
            acts.append(g)
        color = self.color(g)
        if return_activations:
            return density, color, acts
        return density, color

    def forward(self, rays_o, rays_d, t, states_x=None, states_d=None):
        pts, ex, ed = self.encode(rays_o, rays_d, t, states_x, states_d)
        density, color = self.mlp(ex, ed)
        return pts, density, color

    # -- rendering -----------------------------------------------------------
    def render_rays_full(self, rays_o, rays_d, near, far, num_samples, u=None,
                         noise=None, states_x=None, states_d=None, t=None):
        """Returns dict(pixels[B,3], depth[B], acc[B], weights[B,S], t[B,S],
        density[B,S,1], color[B,S,3]).  ``u`` = stratified uniforms or None,
        ``noise`` = N(0,1)*std tensor [B,S,1] already scaled, or None."""
        if t is None:
            t = stratified_t(near, far, num_samples, rays_o.shape[:-1], u=u,
                             dtype=rays_o.dtype, device=rays_o.device)
        pts, density, color = self.forward(rays_o, rays_d, t, states_x, states_d)
        if noise is not None:
            density = density + noise
        w = alpha_compositing_coefficients(pts, density)
        pixels = (w * torch.sigmoid(color)).sum(dim=-2)
        depth = (w[..., 0] * t).sum(dim=-1)
        acc = w[..., 0].sum(dim=-1)
        return dict(pixels=pixels, depth=depth, acc=acc, weights=w[..., 0], t=t,
                    density=density, color=color, points=pts)

    def render_rays(self, rays_o, rays_d, near, far, num_samples, states_x=None,
                    states_d=None, randomly_sample=False, density_noise_std=0.0):
        """Notebook cell 8 signature: -> pixels [B,3].  Draw order of the global
        RNG: rand(u) first, then randn(noise) (HEAD draws randn even when std=0,
        /root/reference/nerf/model.py:652-654)."""
        u = None
        if randomly_sample:
            u = torch.rand(*rays_o.shape[:-1], num_samples,
                           dtype=rays_o.dtype, device=rays_o.device)
        noise = torch.randn(*rays_o.shape[:-1], num_samples, 1,
                            dtype=rays_o.dtype, device=rays_o.device) * density_noise_std
        return self.render_rays_full(rays_o, rays_d, near, far, num_samples, u=u,
                                     noise=noise, states_x=states_x,
                                     states_d=states_d)["pixels"]

    def render_rays_hierarchical(self, rays_o, rays_d, near, far, num_coarse,
                                 num_fine, states_x=None, states_d=None):
        """D3: 64 coarse (deterministic) + num_fine inverse-CDF samples; the fine
        pass evaluates the sorted union (num_coarse + num_fine points) with the
        SAME network.  Returns the fine-pass dict."""
        coarse = self.render_rays_full(rays_o, rays_d, near, far, num_coarse,
                                       states_x=states_x, states_d=states_d)
        t_fine = sample_pdf(coarse["t"], coarse["weights"], num_fine)
        t_all, _ = torch.sort(torch.cat([coarse["t"], t_fine], dim=-1), dim=-1)
        return self.render_rays_full(rays_o, rays_d, near, far, num_coarse + num_fine,
                                     states_x=states_x, states_d=states_d, t=t_all)

    def render_image(self, camera_o, camera_r, image_h, image_w, focal_length,
                     near, far, num_samples, states_x=None, states_d=None,
                     max_chunk_size=1024):
        """/root/reference/nerf/model.py:727-770 with near/far inserted (nb cell 8)."""
        kwargs = dict(dtype=camera_o.dtype, device=camera_o.device)
        bc = camera_o.shape[0]
        rays = generate_rays(image_h, image_w, focal_length, **kwargs)
        rays = torch.broadcast_to(rays.unsqueeze(0), [bc, image_h, image_w, 3])
        co = torch.broadcast_to(camera_o[:, None, None], [bc, image_h, image_w, 3])
        cr = torch.broadcast_to(camera_r[:, None, None], [bc, image_h, image_w, 3, 3])
        rays_o, rays_d = rays_to_world_coordinates(rays, co, cr)
        rays_o = torch.flatten(rays_o, 0, 2)
        rays_d = torch.flatten(rays_d, 0, 2)
        sx = sd = None
        if states_x is not None:
            sx = torch.broadcast_to(states_x[:, None, None], [bc, image_h, image_w, states_x.shape[-1]]).flatten(0, 2)
        if states_d is not None:
            sd = torch.broadcast_to(states_d[:, None, None], [bc, image_h, image_w, states_d.shape[-1]]).flatten(0, 2)
        out = []
        for i in range(0, rays_o.shape[0], max_chunk_size):
            sl = slice(i, i + max_chunk_size)
            out.append(self.render_rays_full(
                rays_o[sl], rays_d[sl], near, far, num_samples,
                states_x=None if sx is None else sx[sl],
                states_d=None if sd is None else sd[sl])["pixels"])
        return torch.cat(out, dim=0).reshape(bc, image_h, image_w, 3)


# -- camera helpers used by the fixtures (tiny-NeRF convention) --
def pose_spherical(theta_deg, phi_deg, radius):
    th, ph = math.radians(theta_deg), math.radians(phi_deg)
    trans = np.eye(4); trans[2, 3] = radius
    rot_phi = np.array([[1, 0, 0, 0], [0, math.cos(ph), -math.sin(ph), 0],
                        [0, math.sin(ph), math.cos(ph), 0], [0, 0, 0, 1]], dtype=np.float64)
    rot_theta = np.array([[math.cos(th), 0, -math.sin(th), 0], [0, 1, 0, 0],
                          [math.sin(th), 0, math.cos(th), 0], [0, 0, 0, 1]], dtype=np.float64)
    flip = np.array([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]], dtype=np.float64)
    return torch.tensor(flip @ rot_theta @ rot_phi @ trans, dtype=torch.float32)


TINY_NERF_FOCAL = 138.88887889922103
FIXTURE_POSE = (180.0, -26.3125, 4.03)
