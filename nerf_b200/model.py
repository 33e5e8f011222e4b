"""Drop-in `NeRF` module: the method surface of the reference's nerf/model.py, with the
bodies dispatching to the sm_100a kernels in libnerf_b200.so.

Generations:
  * G1  -- `NeRF(normalize_position=6.0)`; `render_rays(o, d, near, far, S, ...) -> [B,3]`;
           the model examples/nerf.pth was saved from (example.ipynb cells 5-9).
  * G2  -- `NeRF(normalize_position=, density_inputs=3+Sx)`;
           `render_rays(..., states_x=) -> [B,stages,3]` (train_conditional_nerf.py:103-132).
  * static helpers shared with HEAD: generate_rays, rays_to_world_coordinates,
           alpha_compositing_coefficients, spherical_to_cartesian, get_rotation_matrix
           (nerf/model.py:242-367, 437-469).
"""
import math
import os

import torch
import torch.nn as nn

from . import ops

PE_X_OCTAVES, PE_D_OCTAVES, HIDDEN = 10, 6, 256


def _unit(n_in, n_out):
    return [nn.Linear(n_in, n_out), nn.ReLU(), nn.LayerNorm(n_out)]


def g1_param_layout_py(dx, dd):
    """Pure-python mirror of nerf_g1_param_layout (state_dict order of nerf.pth)."""
    h, ex, ed = HIDDEN, 2 * PE_X_OCTAVES * dx, 2 * PE_D_OCTAVES * dd
    sizes = [h, 1, 3 * h, 3]
    for in0, units in ((ex, 4), (h + ex, 4), (h + ed, 2)):
        for u in range(units):
            sizes += [h * (in0 if u == 0 else h), h, h, h]
    off, p = [], 0
    for s in sizes:
        off.append(p); p += s
    return off, p


class NeRF(nn.Module):

    def __init__(self, normalize_position=1.0, density_inputs=None, color_inputs=None,
                 hidden_size=HIDDEN, precision=None, normalize_direction=True,
                 stages_axis=None, max_chunk_rays=32768):
        super().__init__()
        if hidden_size != HIDDEN:
            raise ValueError("the sm_100a kernels are built for hidden_size=256 (examples/nerf.pth)")
        if precision is None:
            # the reference's call sites (`NeRF(normalize_position=6.0)`, train_conditional_nerf.py:103-104) carry no
            # precision argument: NERF_B200_PRECISION picks the contraction for them (default: the fp32 parity mode)
            precision = os.environ.get("NERF_B200_PRECISION", "fp32")
        if precision not in ops.PRECISIONS:
            raise ValueError(f"precision must be one of {list(ops.PRECISIONS)}")
        # G2 call sites pass density_inputs explicitly and expect a [B,stages,3] result
        self.stages_axis = (density_inputs is not None) if stages_axis is None else bool(stages_axis)
        self.normalize_position = float(normalize_position)
        self.density_inputs = 3 if density_inputs is None else int(density_inputs)
        self.color_inputs = 3 if color_inputs is None else int(color_inputs)
        self.hidden_size = hidden_size
        self.precision = precision
        self.normalize_direction = bool(normalize_direction)
        self.max_chunk_rays = int(max_chunk_rays)
        ex = 2 * PE_X_OCTAVES * self.density_inputs
        ed = 2 * PE_D_OCTAVES * self.color_inputs
        h = hidden_size
        # registration order == key order of examples/nerf.pth
        self.density = nn.Linear(h, 1)
        self.color = nn.Linear(h, 3)
        self.block_0 = nn.Sequential(*(_unit(ex, h) + _unit(h, h) + _unit(h, h) + _unit(h, h)))
        self.block_1 = nn.Sequential(*(_unit(h + ex, h) + _unit(h, h) + _unit(h, h) + _unit(h, h)))
        self.block_2 = nn.Sequential(*(_unit(h + ed, h) + _unit(h, h)))
        self._flat = None
        self._ladders = {}
        self._flatten()

    # ---------------------------------------------------------- flat storage ----
    def _flatten(self):
        """Make the 44 parameters views of one contiguous fp32 vector (state_dict
        order) so the kernels, the gradient all-reduce and Adam see a single buffer."""
        params = list(self.parameters())
        off, total = g1_param_layout_py(self.density_inputs, self.color_inputs)
        assert len(params) == 44 and sum(p.numel() for p in params) == total
        dev = params[0].device
        flat = torch.empty(total, dtype=torch.float32, device=dev)
        with torch.no_grad():
            for p, o in zip(params, off):
                flat[o:o + p.numel()].copy_(p.detach().reshape(-1).float())
                p.data = flat[o:o + p.numel()].view(p.shape)
        self._flat, self._offsets = flat, off
        # where each parameter lives, in parameters() order: the per-step aliasing check reads these 44 slots directly
        # instead of walking the module tree (0.19 ms per call through self.parameters())
        self._slots = [(mod, name) for mod in self.modules() for name, q in mod._parameters.items() if q is not None]
        assert len(self._slots) == len(params) and all(mod._parameters[name] is q for (mod, name), q in zip(self._slots, params))
        return flat

    def _flat_ok(self):
        f = self._flat
        if f is None:
            return False
        base, dev = f.data_ptr(), f.device
        for (mod, name), o in zip(self._slots, self._offsets):
            p = mod._parameters.get(name)                      # (a parameter that was replaced or moved fails the test below)
            if p is None or p.data_ptr() != base + 4 * o or p.device != dev or p.dtype != torch.float32:
                return False
        return True

    def flat_parameters(self):
        if not self._flat_ok():
            self._flatten()
        return self._flat

    def _param_list(self):
        """list(self.parameters()) without the walk over the module tree (0.15-0.19 ms per call): the 44 slots recorded at
        flatten time, read afresh (a replaced Parameter object is seen)"""
        self.flat_parameters()
        return [mod._parameters[name] for mod, name in self._slots]

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        self._flatten()
        return out

    # --------------------------------------------------------- static helpers ----
    @staticmethod
    def generate_rays(image_h, image_w, focal_length, dtype=torch.float32, device="cuda"):
        """Camera-space rays [H,W,3] (reference nerf/model.py:242-278)."""
        dev = torch.device(device)
        cam_o = torch.zeros(1, 3, device=dev)
        cam_r = torch.eye(3, device=dev).unsqueeze(0)
        _, d = ops.generate_rays(cam_o, cam_r, image_h, image_w, focal_length)
        return d.reshape(image_h, image_w, 3).to(dtype)

    @staticmethod
    def rays_to_world_coordinates(rays, camera_o, camera_r):
        """(reference nerf/model.py:336-367) -- a 3x3 product per ray; host-side torch."""
        return camera_o, (camera_r * rays.unsqueeze(-2)).sum(dim=-1)

    @staticmethod
    def spherical_to_cartesian(yaw, elevation):
        """(reference nerf/model.py:280-306)"""
        return torch.stack([torch.cos(yaw) * torch.cos(elevation),
                            torch.sin(yaw) * torch.cos(elevation), torch.sin(elevation)], dim=-1)

    @staticmethod
    def get_rotation_matrix(eye_vector, up_vector):
        """(reference nerf/model.py:308-334)"""
        return torch.stack([torch.linalg.cross(eye_vector, up_vector, dim=-1), up_vector, -eye_vector], dim=-1)

    @staticmethod
    def direction_to_rotation_matrix(directions):
        """G2 helper (train_conditional_nerf.py:86-87), source absent from the reference
        tree; SURVEY A.4(4): camera looks along `directions`, up derived from world +z."""
        eye = directions / torch.linalg.norm(directions, dim=-1, keepdim=True)
        z = torch.zeros_like(eye); z[..., 2] = 1.0
        right = torch.linalg.cross(eye, z, dim=-1)
        degenerate = torch.linalg.norm(right, dim=-1, keepdim=True) < 1e-6
        alt = torch.zeros_like(eye); alt[..., 1] = 1.0
        right = torch.where(degenerate, torch.linalg.cross(eye, alt, dim=-1), right)
        right = right / torch.linalg.norm(right, dim=-1, keepdim=True)
        up = torch.linalg.cross(right, eye, dim=-1)
        return NeRF.get_rotation_matrix(eye, up)

    @staticmethod
    def alpha_compositing_coefficients(points, density_outputs):
        """Weights [...,S,1] (reference nerf/model.py:437-469) via the warp-scan kernel."""
        shape = density_outputs.shape
        s = shape[-2]
        sigma = density_outputs.reshape(-1, s)
        pts = points.reshape(-1, s, 3)
        rgb = torch.zeros(sigma.shape[0], s, 3, device=sigma.device)
        _, _, _, w = ops.composite(sigma, rgb, points=pts)
        return w.reshape(shape)

    # --------------------------------------------------------------- sampling ----
    def _ladder(self, near, far, num_samples, device):
        """torch.linspace's own output, so bins are bit-identical to the reference's."""
        key = (float(near), float(far), int(num_samples), str(device))
        if key not in self._ladders:
            self._ladders[key] = torch.linspace(near, far, num_samples, dtype=torch.float32).to(device)
        return self._ladders[key]

    def sample_along_rays(self, rays_o, rays_d, near, far, num_samples, states_x=None, states_d=None,
                          randomly_sample=True, u=None):
        b = rays_o.shape[0]
        if randomly_sample and u is None:
            u = torch.rand(b, num_samples, dtype=torch.float32, device=rays_o.device)
        return ops.sample_linear(near, far, num_samples, b, rays_o.device, u=u if randomly_sample else None,
                                 t_base=self._ladder(near, far, num_samples, rays_o.device))

    # ---------------------------------------------------------------- network ----
    def forward(self, rays_o, rays_d, samples, states_x=None, states_d=None):
        """-> (points [B,S,3], density [B,S,1], color [B,S,3])."""
        ex, ed, pts = ops.encode_g1(rays_o, rays_d, samples, self.normalize_position, states_x, states_d,
                                    self.normalize_direction, want_points=True)
        b, s = samples.shape
        flat = self.flat_parameters()
        sigma, rgb = ops.g1_mlp(ex.view(b * s, ex.shape[-1]), ed, s, self.density_inputs, self.color_inputs,
                                ops.PREC_FP32, flat, self._param_list())
        return pts, sigma.view(b, s, 1), rgb.view(b, s, 3)

    def _render_chunk(self, rays_o, rays_d, near, far, num_samples, states_x, states_d, u, noise, t=None):
        if t is None:
            t = ops.sample_linear(near, far, num_samples, rays_o.shape[0], rays_o.device, u=u,
                                  t_base=self._ladder(near, far, num_samples, rays_o.device))
        _, sigma, rgb = self.forward(rays_o, rays_d, t, states_x, states_d)
        if noise is not None:
            sigma = sigma + noise
        pixels, depth, acc, weights = ops.composite(sigma.squeeze(-1), rgb, rays_o=rays_o, rays_d=rays_d, t=t)
        return dict(pixels=pixels, depth=depth, acc=acc, weights=weights, t=t,
                    density=sigma, color=rgb)

    def render_rays_full(self, rays_o, rays_d, near, far, num_samples, states_x=None, states_d=None,
                         u=None, noise=None, t=None):
        """All outputs of one pass (dict); `u`/`noise` are host-supplied draws (parity mode)."""
        rays_o = rays_o.float().contiguous(); rays_d = rays_d.float().contiguous()
        b = rays_o.shape[0]
        if self._train_tc_ok(num_samples, states_x, states_d) and b > 0:
            # tcgen05 training path under autograd: `loss.backward()` of the reference loop runs the tensor-core
            # data-/weight-gradient kernels (ops._G1TrainTC); the composite is its own differentiable node
            t, sigma, rgb = ops.g1_train_tc(self.flat_parameters(), self._param_list(), rays_o, rays_d, num_samples,
                                            near, far, self.normalize_position, self.normalize_direction,
                                            t_base=self._ladder(near, far, num_samples, rays_o.device), t_in=t, u=u,
                                            states_x=states_x, x3=self.precision == "bf16x3")
            if noise is not None:
                sigma = sigma + noise.view(b, num_samples)
            pixels, depth, acc, weights = ops.composite(sigma, rgb, rays_o=rays_o, rays_d=rays_d, t=t)
            return dict(pixels=pixels, depth=depth, acc=acc, weights=weights, t=t, density=sigma.unsqueeze(-1), color=rgb)
        if self._fused_ok(num_samples, states_x, states_d) and noise is None:
            # tcgen05 path: one persistent kernel for sample+encode+MLP, then the compositing kernel
            x3 = self.precision == "bf16x3"
            packed = self._packed(x3)
            sl = lambda x, i, j: None if x is None else x[i:j]
            step = max(b, 1) if states_x is None else 1 << 17      # conditional: 2 KB of bias rows per ray in the workspace
            outs = []
            for i in range(0, max(b, 1), step):
                j = min(i + step, b)
                outs.append(ops.g1_render_fused(packed, num_samples, near, far, self.normalize_position, self.normalize_direction,
                                                rays_o=rays_o[i:j], rays_d=rays_d[i:j], t_in=sl(t, i, j), u=sl(u, i, j),
                                                t_base=self._ladder(near, far, num_samples, rays_o.device),
                                                want=("pixels", "depth", "acc", "weights", "t", "sigma", "rgb"),
                                                states_x=sl(states_x, i, j), x3=x3))
            out = outs[0] if len(outs) == 1 else {k: torch.cat([o[k] for o in outs], dim=0) for k in outs[0]}
            out["density"] = out.pop("sigma").unsqueeze(-1); out["color"] = out.pop("rgb")
            return out
        if b <= self.max_chunk_rays:
            return self._render_chunk(rays_o, rays_d, near, far, num_samples, states_x, states_d, u, noise, t)
        outs = []
        sl = lambda x, i: None if x is None else x[i:i + self.max_chunk_rays]
        for i in range(0, b, self.max_chunk_rays):
            outs.append(self._render_chunk(sl(rays_o, i), sl(rays_d, i), near, far, num_samples,
                                           sl(states_x, i), sl(states_d, i), sl(u, i), sl(noise, i), sl(t, i)))
        return {k: torch.cat([o[k] for o in outs], dim=0) for k in outs[0]}

    def _tc_shape_ok(self, num_samples, states_x, states_d):
        sx = 0 if states_x is None else states_x.shape[-1]
        return (self.precision == "bf16" and num_samples >= 32
                and self.density_inputs == 3 + sx and self.color_inputs == 3 and states_d is None)

    def _x3_ok(self, num_samples, states_x, states_d):
        """fp32-grade tensor-core forward (hi/lo operand split, csrc/fused_tc_x3.cu): unconditional model, no graph
        (with a graph: _train_tc_ok -> ops._G1TrainTC(x3=True))"""
        return (self.precision == "bf16x3" and num_samples >= 32 and self.density_inputs == 3 and self.color_inputs == 3
                and states_x is None and states_d is None and not self._wants_grad())

    def _packed(self, x3):
        flat = self.flat_parameters()
        return ops.g1_pack_weights_x3(flat) if x3 else ops.g1_pack_weights(flat, self.density_inputs, 3)

    def _fused_ok(self, num_samples, states_x, states_d):
        """tcgen05 render path (no graph recorded).  bf16: position-side conditional inputs are supported,
        direction-side ones take the fp32 kernels; bf16x3: the unconditional model."""
        if self._x3_ok(num_samples, states_x, states_d):
            return True
        return self._tc_shape_ok(num_samples, states_x, states_d) and not self._wants_grad()

    def _wants_grad(self):
        return torch.is_grad_enabled() and any(p.requires_grad for p in self._param_list())

    def _train_tc_ok(self, num_samples, states_x, states_d):
        """tcgen05 path with a graph (bf16, or fp32-grade bf16x3 for the unconditional model): when a gradient can be asked for."""
        if not self._wants_grad():
            return False
        if self.precision == "bf16x3":
            return (num_samples >= 32 and self.density_inputs == 3 and self.color_inputs == 3
                    and states_x is None and states_d is None)
        return self._tc_shape_ok(num_samples, states_x, states_d)

    def render_rays(self, rays_o, rays_d, near, far, num_samples, states_x=None, states_d=None,
                    randomly_sample=False, density_noise_std=0.0):
        """G1: -> pixels [B,3] (example.ipynb cell 8); G2: -> [B,stages,3]
        (train_conditional_nerf.py:120-132).  RNG order follows the reference: rand for
        the jitter, then randn for the density noise (drawn even when std == 0,
        nerf/model.py:652-654)."""
        b = rays_o.shape[0]
        dev = rays_o.device
        u = torch.rand(b, num_samples, dtype=torch.float32, device=dev) if randomly_sample else None
        noise = torch.randn(b, num_samples, 1, dtype=torch.float32, device=dev) * density_noise_std
        out = self.render_rays_full(rays_o, rays_d, near, far, num_samples, states_x, states_d,
                                    u=u, noise=noise if density_noise_std != 0.0 else None)
        pixels = out["pixels"]
        return pixels.unsqueeze(-2) if self.stages_axis else pixels

    def render_rays_hierarchical(self, rays_o, rays_d, near, far, num_coarse, num_fine,
                                 states_x=None, states_d=None):
        """64 coarse + num_fine inverse-CDF samples, fine pass over the sorted union (D3)."""
        with torch.no_grad():
            coarse = self.render_rays_full(rays_o, rays_d, near, far, num_coarse, states_x, states_d)
            t_all = ops.sample_pdf_merge(coarse["t"], coarse["weights"], num_fine)
        return self.render_rays_full(rays_o, rays_d, near, far, num_coarse + num_fine, states_x, states_d, t=t_all)

    @staticmethod
    def _launch_pieces(ray_begin, ray_end, total, chunk):
        """How render_image cuts the ray ids [ray_begin, ray_end) of a frame of `total` rays into launches of about `chunk`
        rays: yields (lo, hi, i, j) = launch the rays [lo, hi), keep [i, j).

        A ray's pixel does not depend on the other rays of its launch, with one exception: the composite picks its kernel by
        the launch's ray count (warp-per-ray below ops.COMPOSITE_WARP_BELOW, thread-per-ray from it on; they sum in different
        orders).  A shard is bit-identical to the slice of the frame, at any world size, if each of its rays sees the kernel
        it sees in the frame.  The frame is cut into `chunk`-ray launches; when chunk >= threshold all of them take the
        thread-per-ray kernel except, possibly, a short last one (the frame's tail).  So outside the tail any launch of at
        least `threshold` rays will do: the shard is cut into equal pieces of its own, none below the threshold (a shard
        shorter than the threshold is widened to it and sliced); the tail is launched as it is.  (Aligning the shard's
        launches to the frame's chunks, as this did at first, made a rank with two short edge pieces render 131 072 rays
        for its 80 000 at N = 8.)"""
        thr = ops.COMPOSITE_WARP_BELOW
        ray_end = min(ray_end, total)
        if chunk < thr:                                      # every launch of the frame is below the threshold: so is every piece
            for i in range(ray_begin, ray_end, chunk):
                j = min(i + chunk, ray_end)
                yield i, j, i, j
            return
        last = total % chunk
        tail0 = total - last if 0 < last < thr else total    # rays from tail0 on see the warp-per-ray kernel in the frame
        a, b = ray_begin, min(ray_end, tail0)
        n = b - a
        if 0 < n < thr:                                      # (tail0 >= chunk >= thr here: there is room to widen)
            lo = max(0, min(a, tail0 - thr))
            yield lo, lo + thr, a, b
        elif n > 0:
            k = max(1, min(-(-n // chunk), n // thr))
            for q in range(k):
                i, j = a + n * q // k, a + n * (q + 1) // k
                yield i, j, i, j
        a = max(ray_begin, tail0)
        if a < ray_end:
            yield a, ray_end, a, ray_end

    def render_image(self, camera_o, camera_r, image_h, image_w, focal_length, near, far, num_samples,
                     states_x=None, states_d=None, max_chunk_size=None, num_fine=0,
                     ray_begin=0, ray_end=None):
        """-> [Bc,H,W,3] (example.ipynb cell 8 lines 29-37; flatten order and chunking
        of nerf/model.py:748-770).  `ray_begin/ray_end` render a slice of the flattened
        ray ids (multi-GPU sharding); then the result is [n,3]."""
        bc = camera_o.shape[0]
        total = bc * image_h * image_w
        full = ray_begin == 0 and ray_end is None
        ray_end = total if ray_end is None else ray_end
        chunk = self.max_chunk_rays if max_chunk_size is None else int(max_chunk_size)
        outs = []
        if self._fused_ok(num_samples, states_x, states_d) and num_fine == 0:
            x3 = self.precision == "bf16x3"
            packed = self._packed(x3)
            if states_x is not None:                        # one state row per camera (train_conditional_nerf.py:141-153)
                states_x = states_x.float().reshape(-1, states_x.shape[-1]).expand(bc, -1).contiguous()
            big = 1 << 20                                   # rays per launch (bounds the t/sigma/rgb workspace)
            for lo, hi, i, j in self._launch_pieces(ray_begin, ray_end, total, big):
                r = ops.g1_render_fused(packed, num_samples, near, far, self.normalize_position,
                                        self.normalize_direction,
                                        camera=(camera_o, camera_r, image_h, image_w, focal_length),
                                        ray_begin=lo, num_rays=hi - lo,
                                        t_base=self._ladder(near, far, num_samples, camera_o.device), want=("pixels",),
                                        states_x=states_x, rays_per_state=image_h * image_w, x3=x3)
                outs.append(r["pixels"][i - lo:j - lo])
            if not outs:                                    # an empty shard (more ranks than rays)
                return torch.empty(0, 3, device=camera_o.device)
            img = torch.cat(outs, dim=0)
            return img.reshape(bc, image_h, image_w, 3) if full else img
        for lo, hi, i, j in self._launch_pieces(ray_begin, ray_end, total, chunk):
            o, d = ops.generate_rays(camera_o, camera_r, image_h, image_w, focal_length, lo, hi)
            sx = sd = None
            if states_x is not None or states_d is not None:
                cam = torch.arange(lo, hi, device=o.device) // (image_h * image_w)
                sx = None if states_x is None else states_x.float()[cam].contiguous()
                sd = None if states_d is None else states_d.float()[cam].contiguous()
            if num_fine > 0:
                r = self.render_rays_hierarchical(o, d, near, far, num_samples, num_fine, sx, sd)
            else:
                r = self.render_rays_full(o, d, near, far, num_samples, sx, sd)
            outs.append(r["pixels"][i - lo:j - lo])
        if not outs:
            return torch.empty(0, 3, device=camera_o.device)
        img = torch.cat(outs, dim=0)
        return img.reshape(bc, image_h, image_w, 3) if full else img
