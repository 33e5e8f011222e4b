"""HEAD generation of the reference model -- `nerf/model.py` as it is in the reference tree: mip-NeRF
style integrated positional encoding, a Linear->LayerNorm->ReLU trunk and a 1+3+50 output head with a
segmentation composite.  Same constructor, method names, argument meaning, return shapes and
state_dict keys as the reference class (nerf/model.py:471-770); every arithmetic stage is a kernel of
libnerf_b200.so (csrc/head.cu, csrc/rays.cu).  The reference ships no training entry point for this
generation (train_conditional_nerf.py and the notebook drive the G1/G2 API), but its module is
differentiable: in fp32 mode `render_rays` called with autograd enabled records ONE autograd node
(`_HeadRenderFn`) whose backward is the composite + trunk backward kernels of csrc/head.cu, so
`loss.backward()` fills `.grad` of the 22 prediction_heads tensors exactly as it does on the reference.
"""
import ctypes
import math

import torch
import torch.nn as nn

from . import ops
from ._lib import check, lib, ptr, stream
from .model import NeRF as _G1

LOG2_NEAR = -9.43633744014          # nerf/model.py:414


LDO = 64        # row stride of the output-row gradients (n_out <= 64, zero padded): 16-byte loads in the tiles


class _HeadRenderFn(torch.autograd.Function):
    """render_rays of the HEAD generation (nerf/model.py:596-668) as one autograd node over the 22
    prediction_heads tensors: forward = encoder + trunk (activations kept) + composite, backward =
    nerf_head_composite_bwd + nerf_head_mlp_bwd.  Rays and fenceposts get no gradient (they depend on no
    parameter)."""

    @staticmethod
    def forward(ctx, model, o, d, t, noise, noise_std, *params):
        b, s = t.shape
        s1 = s - 1
        n_seg = model.segmentation_outputs
        n_out = 1 + model.color_outputs + n_seg
        if n_out > LDO:
            raise ValueError("HEAD training supports at most 60 segmentation outputs")
        mean, _, feat = model.integrated_pe(o, d, t)
        feat = feat.view(b * s1, -1)
        ptrs, keep = model._param_pointers()
        nbytes = lib().nerf_head_mlp_train_workspace_bytes(b * s1)
        ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=t.device)
        out = torch.empty(b * s1, n_out, device=t.device)
        check(lib().nerf_head_mlp_fwd_train(ptrs, ptr(feat), b * s1, feat.shape[1], n_out, ptr(out), ptr(ws), nbytes, stream()),
              "nerf_head_mlp_fwd_train")
        img = torch.empty(b, 3, device=t.device); seg = torch.empty(b, n_seg, device=t.device)
        mean = mean.view(b, s1, 3).contiguous()
        check(lib().nerf_head_composite(ptr(out), n_out, n_seg, ptr(mean), ptr(noise), float(noise_std), b, s1,
                                        ptr(img), ptr(seg), None, 0, stream()), "nerf_head_composite")
        ctx.model, ctx.noise_std, ctx.shape = model, float(noise_std), (b, s1, n_out, n_seg)
        ctx.noise = noise
        ctx.save_for_backward(out, mean, feat, ws, seg, *params)
        return img, seg

    @staticmethod
    def backward(ctx, d_img, d_seg):
        out, mean, feat, ws, seg, *params = ctx.saved_tensors
        b, s1, n_out, n_seg = ctx.shape
        d_img = ops._f32c(d_img) if d_img is not None else None
        d_seg = ops._f32c(d_seg) if (d_seg is not None and n_seg > 0) else None
        d_out = torch.empty(b * s1, LDO, device=out.device)
        check(lib().nerf_head_composite_bwd(ptr(out), n_out, n_seg, ptr(mean), ptr(ctx.noise), ctx.noise_std, b, s1,
                                            ptr(seg), ptr(d_img), ptr(d_seg), ptr(d_out), LDO, stream()),
              "nerf_head_composite_bwd")
        sizes = [p.numel() for p in params]
        flat = torch.zeros(sum((n + 3) // 4 * 4 for n in sizes), device=out.device)      # 16-byte aligned views
        grads, off = [], 0
        for p, n in zip(params, sizes):
            grads.append(flat[off:off + n].view_as(p)); off += (n + 3) // 4 * 4
        pp = (ctypes.c_void_p * 22)(*[p.data_ptr() for p in params])
        gp = (ctypes.c_void_p * 22)(*[g.data_ptr() for g in grads])
        check(lib().nerf_head_mlp_bwd(pp, ptr(feat), b * s1, feat.shape[1], n_out, ptr(d_out), LDO, gp, ptr(ws),
                                      ws.numel(), stream()), "nerf_head_mlp_bwd")
        return (None,) * 6 + tuple(grads)


class HeadNeRF(nn.Module):

    def __init__(self, color_outputs=3, segmentation_outputs=50, hidden_size=256, encoding_size=32,
                 focal_length=112.0, min_x=-20.0, min_y=-20.0, min_z=-20.0,
                 max_x=20.0, max_y=20.0, max_z=20.0, max_chunk_rays=65536, precision="fp32"):
        super().__init__()
        if precision not in ("fp32", "bf16"):
            raise ValueError("precision must be 'fp32' or 'bf16'")
        if precision == "bf16" and (encoding_size != 32 or 1 + color_outputs + segmentation_outputs > 64):
            raise ValueError("the tcgen05 HEAD kernel is built for encoding_size=32 and at most 64 outputs")
        self.precision = precision
        if hidden_size != 256:
            raise ValueError("the sm_100a kernels are built for hidden_size=256")
        if color_outputs != 3 or not (0 <= segmentation_outputs <= 64):
            raise ValueError("kernels support color_outputs=3 and segmentation_outputs<=64")
        if encoding_size % 2 or not (2 <= encoding_size <= 64):
            raise ValueError("encoding_size must be even and <= 64")
        self.color_outputs, self.segmentation_outputs = color_outputs, segmentation_outputs
        self.hidden_size, self.encoding_size = hidden_size, encoding_size
        self.focal_length = focal_length
        self.max_chunk_rays = int(max_chunk_rays)
        # nerf/model.py:519-523: the world box is a registered buffer, hence part of state_dict
        self.register_buffer("rays_min", torch.tensor([[[min_x, min_y, min_z]]], dtype=torch.float32))
        self.register_buffer("rays_max", torch.tensor([[[max_x, max_y, max_z]]], dtype=torch.float32))
        layers, n_in = [], 3 * encoding_size
        for _ in range(5):                                      # nerf/model.py:525-540
            layers += [nn.Linear(n_in, hidden_size), nn.LayerNorm(hidden_size), nn.ReLU()]
            n_in = hidden_size
        layers.append(nn.Linear(hidden_size, 1 + color_outputs + segmentation_outputs))
        self.prediction_heads = nn.Sequential(*layers)
        self._ladders = {}

    # shared static helpers (same kernels as the G1 class)
    generate_rays = staticmethod(_G1.generate_rays)
    rays_to_world_coordinates = staticmethod(_G1.rays_to_world_coordinates)
    spherical_to_cartesian = staticmethod(_G1.spherical_to_cartesian)
    get_rotation_matrix = staticmethod(_G1.get_rotation_matrix)
    alpha_compositing_coefficients = staticmethod(_G1.alpha_compositing_coefficients)

    # ------------------------------------------------------------------ pieces ----
    def _param_pointers(self):
        ps = [p.detach() for p in self.prediction_heads.parameters()]
        assert len(ps) == 22
        for p in ps:
            if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                raise RuntimeError("HEAD kernels need contiguous fp32 CUDA parameters (there is no CPU path)")
        return (ctypes.c_void_p * 22)(*[p.data_ptr() for p in ps]), ps

    def _box_diag(self):
        return float(torch.linalg.norm((self.rays_max - self.rays_min).float()).item())

    def _ladder(self, num_samples, device):
        key = (num_samples, str(device))
        if key not in self._ladders:                             # torch's own linspace/pow -> bit-identical bins
            self._ladders[key] = torch.pow(2.0, torch.linspace(LOG2_NEAR, 0.0, num_samples, dtype=torch.float32)).to(device)
        return self._ladders[key]

    def sample_along_rays(self, rays_o, rays_d, num_samples, states_x=None, states_d=None,
                          randomly_sample=True, u=None, seed=None, ray_id_base=0):
        """nerf/model.py:369-435: geometric ladder x world-box diagonal, stratified jitter.  The jitter is a fresh
        `torch.rand` draw per call like the reference's (:432, so `torch.manual_seed` governs it), unless `u` is
        handed in or `seed` selects the counter-based stream keyed by (seed, ray_id_base + ray, sample)."""
        batch = rays_o.shape[:-1]
        b = int(math.prod(batch)) if len(batch) else 1
        if not randomly_sample:
            u = None
        elif u is None and seed is None:
            u = torch.rand(b, num_samples, dtype=torch.float32, device=rays_o.device)
        t = ops.sample_geometric(self._box_diag(), num_samples, b, rays_o.device, u=u,
                                 jitter=bool(randomly_sample) and u is None, seed=0 if seed is None else seed,
                                 ray_id_base=ray_id_base, t_base=self._ladder(num_samples, rays_o.device))
        return t.view(*batch, num_samples)

    def integrated_pe(self, rays_o, rays_d, samples):
        """nerf/model.py:544-551 -> (means [..,S-1,3], covs [..,S-1,3], features [..,S-1,3*encoding_size])."""
        batch = samples.shape[:-1]; s = samples.shape[-1]
        o = ops._f32c(rays_o.reshape(-1, 3)); d = ops._f32c(rays_d.reshape(-1, 3)); t = ops._f32c(samples.reshape(-1, s))
        b = t.shape[0]; n_deg = self.encoding_size // 2
        mean = torch.empty(b, s - 1, 3, device=t.device); cov = torch.empty(b, s - 1, 3, device=t.device)
        feat = torch.empty(b, s - 1, 6 * n_deg, device=t.device)
        r = 1.0 / (math.sqrt(3) * self.focal_length)             # :546 -- the CONSTRUCTOR focal
        check(lib().nerf_head_encode(ptr(o), ptr(d), ptr(t), b, s, r, -4, n_deg, ptr(mean), ptr(cov), ptr(feat), stream()),
              "nerf_head_encode")
        return (mean.view(*batch, s - 1, 3), cov.view(*batch, s - 1, 3), feat.view(*batch, s - 1, 6 * n_deg))

    def _fwd_tc(self, o, d, t, log_softmax_seg=False):
        """Encoder + trunk + output head on tcgen05 (csrc/head_tc.cu) -> (means [B,S-1,3], out [B,S-1,n_out])."""
        b, s = t.shape
        n_out = 1 + self.color_outputs + self.segmentation_outputs
        ptrs, keep = self._param_pointers()
        packed = torch.empty(lib().nerf_head_packed_bytes(), dtype=torch.uint8, device=t.device)
        check(lib().nerf_head_pack_weights(ptrs, 3 * self.encoding_size, n_out, ptr(packed), stream()), "nerf_head_pack_weights")
        out = torch.empty(b, s - 1, n_out, device=t.device); mean = torch.empty(b, s - 1, 3, device=t.device)
        r = 1.0 / (math.sqrt(3) * self.focal_length)
        check(lib().nerf_head_fwd_tc(ptr(packed), ptr(o), ptr(d), ptr(t), b, s, r, -4, self.encoding_size // 2, n_out,
                                     int(log_softmax_seg), ptr(out), ptr(mean), stream()), "nerf_head_fwd_tc")
        return mean, out

    def _trunk(self, feat2d):
        p, n_in = feat2d.shape
        n_out = 1 + self.color_outputs + self.segmentation_outputs
        out = torch.empty(p, n_out, device=feat2d.device)
        nbytes = lib().nerf_head_mlp_workspace_bytes(p)
        ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=feat2d.device)
        ptrs, keep = self._param_pointers()
        check(lib().nerf_head_mlp_fwd(ptrs, ptr(feat2d), p, n_in, n_out, ptr(out), ptr(ws), nbytes, stream()),
              "nerf_head_mlp_fwd")
        return out

    @torch.no_grad()
    def forward(self, rays_o, rays_d, samples, states_x=None, states_d=None):
        """nerf/model.py:553-594 -> (mean, density [..,1], color [..,3], segmentation [..,n_seg]).
        Inference only: the differentiable entry point of this generation is `render_rays` (one autograd node
        over encoder + trunk + composite); the raw network outputs returned here carry no graph."""
        if self.precision == "bf16":
            batch = samples.shape[:-1]; s = samples.shape[-1]
            mean, out = self._fwd_tc(ops._f32c(rays_o.reshape(-1, 3)), ops._f32c(rays_d.reshape(-1, 3)), ops._f32c(samples.reshape(-1, s)))
            mean, out = mean.view(*batch, s - 1, 3), out.view(*batch, s - 1, -1)
        else:
            mean, _, feat = self.integrated_pe(rays_o, rays_d, samples)
            out = self._trunk(feat.reshape(-1, feat.shape[-1])).view(*feat.shape[:-1], -1)
        density, color, seg = out.split([1, self.color_outputs, self.segmentation_outputs], dim=-1)
        return mean, density, color, seg

    def render_rays(self, rays_o, rays_d, num_samples, states_x=None, states_d=None,
                    randomly_sample=False, density_noise_std=0.0, u=None, noise=None, seed=None, ray_id_base=0):
        """nerf/model.py:596-668 -> (image [B,1,3], segmentation [B,1,n_seg]).  With autograd enabled and
        parameters that require grad (fp32 mode) the result carries the graph of `_HeadRenderFn`; the
        tcgen05 (bf16) form is inference only.  RNG like the reference: `torch.rand` for the jitter (:432), then
        `torch.randn` for the density noise, drawn even when std == 0 (:652-654)."""
        params = list(self.prediction_heads.parameters())
        if torch.is_grad_enabled() and self.precision == "fp32" and any(p.requires_grad for p in params):
            with torch.no_grad():
                o = ops._f32c(rays_o.reshape(-1, 3)); d = ops._f32c(rays_d.reshape(-1, 3))
                b, s1 = o.shape[0], num_samples - 1
                t = self.sample_along_rays(o, d, num_samples, randomly_sample=randomly_sample, u=u, seed=seed,
                                           ray_id_base=ray_id_base)
                if noise is None:                                # the reference draws torch.randn here, whatever std (:652)
                    noise = torch.randn(b, s1, device=o.device)
                nz = ops._f32c(noise.reshape(b, s1)) if density_noise_std != 0.0 else None
            img, seg = _HeadRenderFn.apply(self, o, d, t, nz, float(density_noise_std), *params)
            return img.unsqueeze(-2), seg.unsqueeze(-2)
        with torch.no_grad():
            return self._render_rays_eval(rays_o, rays_d, num_samples, randomly_sample, density_noise_std, u, noise, seed,
                                          ray_id_base)

    def _render_rays_eval(self, rays_o, rays_d, num_samples, randomly_sample, density_noise_std, u, noise, seed, ray_id_base=0):
        o = ops._f32c(rays_o.reshape(-1, 3)); d = ops._f32c(rays_d.reshape(-1, 3))
        b = o.shape[0]
        t = self.sample_along_rays(o, d, num_samples, randomly_sample=randomly_sample, u=u, seed=seed, ray_id_base=ray_id_base)
        s1 = num_samples - 1
        if self.precision == "bf16":
            mean, out = self._fwd_tc(o, d, t, log_softmax_seg=True)      # :662 fused into the output head's epilogue
            out = out.view(b * s1, -1)
        else:
            mean, _, feat = self.integrated_pe(o, d, t)
            out = self._trunk(feat.view(b * s1, -1))
        if noise is None:                                        # the reference draws torch.randn here, whatever std (:652)
            noise = torch.randn(b, s1, device=o.device)
        img = torch.empty(b, 3, device=o.device); seg = torch.empty(b, self.segmentation_outputs, device=o.device)
        nz = ops._f32c(noise.reshape(b, s1)) if density_noise_std != 0.0 else None
        check(lib().nerf_head_composite(ptr(out), out.shape[1], self.segmentation_outputs, ptr(mean.view(b, s1, 3).contiguous()),
                                        ptr(nz), float(density_noise_std), b, s1, ptr(img), ptr(seg), None,
                                        int(self.precision == "bf16"), stream()),
              "nerf_head_composite")
        return img.unsqueeze(-2), seg.unsqueeze(-2)

    @torch.no_grad()
    def render_image(self, camera_o, camera_r, image_h, image_w, focal_length, num_samples,
                     states_x=None, states_d=None, max_chunk_size=1024, randomly_sample=False,
                     density_noise_std=0.0, ray_begin=0, ray_end=None, seed=None):
        """nerf/model.py:670-770 -> (image [Bc,H,W,3], segmentation [Bc,H,W,n_seg]).  Rays are generated
        on the device from their flattened id; chunks are independent (max_chunk_size only bounds
        memory here, so it is raised to max_chunk_rays).  ray_begin/ray_end render a shard.  A jittered frame
        uses ONE seed (drawn from torch's generator when not given) with the counter-based stream keyed by the
        frame's ray ids, so chunk sizes and shard boundaries do not change the draw."""
        camera_o, camera_r = ops._f32c(camera_o), ops._f32c(camera_r)
        bc = camera_o.shape[0]
        total = bc * image_h * image_w
        ray_end = total if ray_end is None else ray_end
        chunk = max(int(max_chunk_size), self.max_chunk_rays)
        if seed is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item()) if randomly_sample else 0
        imgs, segs = [], []
        for r0 in range(ray_begin, ray_end, chunk):
            r1 = min(r0 + chunk, ray_end)
            o, d = ops.generate_rays(camera_o, camera_r, image_h, image_w, focal_length, r0, r1)
            a, s = self.render_rays(o, d, num_samples, randomly_sample=randomly_sample,
                                    density_noise_std=density_noise_std, seed=seed, ray_id_base=r0)
            imgs.append(a[:, -1]); segs.append(s[:, -1])
        img, seg = torch.cat(imgs), torch.cat(segs)
        if ray_begin == 0 and ray_end == total:
            return (img.view(bc, image_h, image_w, self.color_outputs),
                    seg.view(bc, image_h, image_w, self.segmentation_outputs))
        return img, seg
