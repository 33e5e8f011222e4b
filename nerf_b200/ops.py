"""Host-side operators over the C ABI (include/nerf_b200.h).  PyTorch supplies device
memory, streams and autograd plumbing only; all arithmetic is in libnerf_b200.so."""
import atexit
import collections
import ctypes
import json
import os

import torch

from . import _lib
from ._lib import check, lib, ptr, stream

# precision of the MLP contraction: "fp32" = FFMA tiles (the <= 1e-3 parity mode, forward and backward); "bf16" = tcgen05
# kind::f16 (forward, backward, training); "bf16x3" = fp32-grade operands on tcgen05 (hi/lo split, three products per term):
# forward / render, and training (forward, data-gradient chain and weight-gradient GEMMs) of the unconditional model
PREC_FP32, PREC_BF16, PREC_BF16X3 = 0, 1, 2
PRECISIONS = {"fp32": PREC_FP32, "bf16": PREC_BF16, "bf16x3": PREC_BF16X3}
HIDDEN = 256


# nerf_composite_fwd takes the warp-per-ray kernels below this many rays and the thread-per-ray tile kernels from it on
# (csrc/composite.cu); the two sum in different orders
COMPOSITE_WARP_BELOW = 1 << 14

# Calls per kernel family made by this process (which path ran is a testable fact): NERF_B200_STATS=<file> dumps
# them as JSON at exit, so a test that runs an unmodified reference entry point in a subprocess can assert that
# its `render_rays / loss.backward()` loop reached the tcgen05 kernels.
COUNTERS = collections.Counter()


def _dump_counters():
    path = os.environ.get("NERF_B200_STATS")
    if path:
        with open(path, "w") as f:
            json.dump(dict(COUNTERS), f)


atexit.register(_dump_counters)


def _f32c(t):
    if t is None:
        return None
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


# ------------------------------------------------------------------ rays -------
def generate_rays(camera_o, camera_r, image_h, image_w, focal_length, ray_begin=0, ray_end=None):
    """World-space rays of the flattened ids [ray_begin, ray_end) of render_image
    (reference nerf/model.py:727-753)."""
    camera_o, camera_r = _f32c(camera_o), _f32c(camera_r)
    n_cams = camera_o.shape[0]
    total = n_cams * image_h * image_w
    ray_end = total if ray_end is None else ray_end
    n = ray_end - ray_begin
    o = torch.empty(n, 3, device=camera_o.device, dtype=torch.float32)
    d = torch.empty(n, 3, device=camera_o.device, dtype=torch.float32)
    check(lib().nerf_generate_rays(ptr(camera_o), ptr(camera_r), n_cams, image_h, image_w,
                                   float(focal_length), ray_begin, ray_end, ptr(o), ptr(d), stream()),
          "nerf_generate_rays")
    return o, d


def gather_pixel_rays(images, poses, focal_length, idx, states_x=None):
    """Batched PixelRayDataset.__getitem__ (reference nerf/dataset.py:283-316)."""
    images, poses = _f32c(images), _f32c(poses)
    idx = idx.to(device=images.device, dtype=torch.int64).contiguous()
    n, h, w, c = images.shape
    b = idx.numel()
    dev = images.device
    o = torch.empty(b, 3, device=dev); d = torch.empty(b, 3, device=dev)
    px = torch.empty(b, c, device=dev)
    bhw = torch.empty(b, 3, device=dev, dtype=torch.int64)
    sx = 0 if states_x is None else states_x.shape[-1]
    st = None
    if sx:
        states_x = _f32c(states_x)
        st = torch.empty(b, sx, device=dev)
    check(lib().nerf_gather_pixel_rays(ptr(images), ptr(poses), poses.shape[1], ptr(states_x) if sx else None, sx,
                                       n, h, w, c, float(focal_length), ptr(idx), b, ptr(o), ptr(d), ptr(px),
                                       ptr(st), ptr(bhw), stream()), "nerf_gather_pixel_rays")
    return o, d, px, st, bhw


def sample_linear(near, far, num_samples, num_rays, device, u=None, jitter=False, seed=0,
                  ray_id_base=0, t_base=None):
    t = torch.empty(num_rays, num_samples, device=device, dtype=torch.float32)
    u = _f32c(u)
    check(lib().nerf_sample_linear(float(near), float(far), ptr(t_base), num_samples, num_rays, ptr(u),
                                   int(bool(jitter) or u is not None), seed, ray_id_base, ptr(t), stream()),
          "nerf_sample_linear")
    return t


def sample_geometric(box_diag, num_samples, num_rays, device, u=None, jitter=False, seed=0,
                     ray_id_base=0, t_base=None):
    t = torch.empty(num_rays, num_samples, device=device, dtype=torch.float32)
    u = _f32c(u)
    check(lib().nerf_sample_geometric(float(box_diag), ptr(t_base), num_samples, num_rays, ptr(u),
                                      int(bool(jitter) or u is not None), seed, ray_id_base, ptr(t), stream()),
          "nerf_sample_geometric")
    return t


def add_density_noise_(sigma, std, seed=0, ray_id_base=0):
    """In place sigma [B,S] += N(0,1)*std with the counter-based generator."""
    b, s = sigma.shape
    check(lib().nerf_add_density_noise(ptr(sigma), s, b, float(std), seed, ray_id_base, stream()),
          "nerf_add_density_noise")
    return sigma


def sample_pdf_merge(t, weights, num_fine, u_fine=None):
    t, weights, u_fine = _f32c(t), _f32c(weights.detach()), _f32c(u_fine)
    b, s = t.shape
    out = torch.empty(b, s + num_fine, device=t.device, dtype=torch.float32)
    check(lib().nerf_sample_pdf_merge(ptr(t), ptr(weights), s, num_fine, ptr(u_fine), b, ptr(out), stream()),
          "nerf_sample_pdf_merge")
    return out


def encode_g1(rays_o, rays_d, t, normalize_position, states_x=None, states_d=None,
              normalize_direction=True, want_points=False):
    rays_o, rays_d, t = _f32c(rays_o), _f32c(rays_d), _f32c(t)
    b, s = t.shape
    sx = 0 if states_x is None else states_x.shape[-1]
    sd = 0 if states_d is None else states_d.shape[-1]
    states_x = _f32c(states_x) if sx else None
    states_d = _f32c(states_d) if sd else None
    dev = t.device
    ex = torch.empty(b, s, 20 * (3 + sx), device=dev)
    ed = torch.empty(b, 12 * (3 + sd), device=dev)
    pts = torch.empty(b, s, 3, device=dev) if want_points else None
    check(lib().nerf_encode_g1(ptr(rays_o), ptr(rays_d), ptr(t), ptr(states_x), sx, ptr(states_d), sd,
                               b, s, float(normalize_position), int(normalize_direction),
                               ptr(ex), ptr(ed), ptr(pts), stream()), "nerf_encode_g1")
    return ex, ed, pts


# ------------------------------------------------------------ compositing ------
class _Composite(torch.autograd.Function):
    """pixels/depth/acc/weights = composite(sigma, rgb | geometry); reference
    nerf/model.py:462-469, 660.  Geometry carries no gradient."""

    @staticmethod
    def forward(ctx, sigma, rgb, points, rays_o, rays_d, t):
        sigma, rgb = _f32c(sigma), _f32c(rgb)
        b, s = sigma.shape
        dev = sigma.device
        pixels = torch.empty(b, 3, device=dev); depth = torch.empty(b, device=dev) if t is not None else None
        acc = torch.empty(b, device=dev); weights = torch.empty(b, s, device=dev)
        check(lib().nerf_composite_fwd(ptr(sigma), ptr(rgb), ptr(points), ptr(rays_o), ptr(rays_d), ptr(t),
                                       b, s, ptr(pixels), ptr(depth), ptr(acc), ptr(weights), stream()),
              "nerf_composite_fwd")
        ctx.save_for_backward(sigma, rgb, points, rays_o, rays_d, t, pixels, depth, acc)
        if depth is None:
            depth = torch.zeros(b, device=dev)
        return pixels, depth, acc, weights

    @staticmethod
    def backward(ctx, g_pixels, g_depth, g_acc, g_weights):
        sigma, rgb, points, rays_o, rays_d, t, pixels, depth, acc = ctx.saved_tensors
        b, s = sigma.shape
        d_sigma = torch.empty_like(sigma); d_rgb = torch.empty_like(rgb)
        g_pixels, g_acc, g_weights = _f32c(g_pixels), _f32c(g_acc), _f32c(g_weights)
        g_depth = _f32c(g_depth) if t is not None else None
        check(lib().nerf_composite_bwd(ptr(sigma), ptr(rgb), ptr(points), ptr(rays_o), ptr(rays_d), ptr(t),
                                       b, s, ptr(g_pixels), ptr(g_depth), ptr(g_acc), ptr(g_weights),
                                       ptr(pixels), ptr(depth), ptr(acc),
                                       ptr(d_sigma), ptr(d_rgb), stream()), "nerf_composite_bwd")
        return d_sigma, d_rgb, None, None, None, None


def composite(sigma, rgb, points=None, rays_o=None, rays_d=None, t=None):
    """sigma [B,S], rgb [B,S,3]; positions from points [B,S,3] or (rays_o, rays_d, t).
    Returns pixels [B,3], depth [B], acc [B], weights [B,S]."""
    return _Composite.apply(sigma, rgb, _f32c(points), _f32c(rays_o), _f32c(rays_d), _f32c(t))


# --------------------------------------------------------------------- MLP -----
def g1_param_layout(dx, dd):
    off = (ctypes.c_int64 * 44)()
    total = lib().nerf_g1_param_layout(dx, dd, off)
    if total < 0:
        raise ValueError("bad Dx/Dd")
    return list(off), int(total)


class _G1MLP(torch.autograd.Function):
    """sigma [P], rgb [P,3] = MLP(ex [P,20Dx], ed [P/S,12Dd]); SURVEY A.3.  `flat` is
    the flat fp32 parameter vector the 44 `params` alias; grads are produced in one
    flat buffer and handed back as 44 views (one all-reduce / one Adam launch)."""

    @staticmethod
    def forward(ctx, ex, ed, samples_per_ray, dx, dd, precision, flat, *params):
        p = ex.shape[0]
        need_grad = any(ctx.needs_input_grad)
        nbytes = lib().nerf_g1_mlp_workspace_bytes(p, dx, dd, int(need_grad))
        ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=ex.device)
        sigma = torch.empty(p, device=ex.device); rgb = torch.empty(p, 3, device=ex.device)
        check(lib().nerf_g1_mlp_fwd(ptr(flat), ptr(ex), ptr(ed), p, samples_per_ray, dx, dd,
                                    int(need_grad), ptr(sigma), ptr(rgb), ptr(ws), nbytes, stream()),
              "nerf_g1_mlp_fwd")
        COUNTERS["g1_mlp_fwd_fp32"] += 1
        ctx.save_for_backward(ex, ed, flat, ws)
        ctx.cfg = (samples_per_ray, dx, dd, precision, [tuple(q.shape) for q in params], nbytes)
        return sigma, rgb

    @staticmethod
    def backward(ctx, d_sigma, d_rgb):
        ex, ed, flat, ws = ctx.saved_tensors
        s, dx, dd, precision, shapes, nbytes = ctx.cfg
        p = ex.shape[0]
        grads = torch.zeros_like(flat)
        d_sigma = _f32c(d_sigma) if d_sigma is not None else torch.zeros(p, device=ex.device)
        d_rgb = _f32c(d_rgb) if d_rgb is not None else torch.zeros(p, 3, device=ex.device)
        check(lib().nerf_g1_mlp_bwd(ptr(flat), ptr(ex), ptr(ed), p, s, dx, dd, ptr(d_sigma),
                                    ptr(d_rgb), ptr(grads), ptr(ws), nbytes, stream()), "nerf_g1_mlp_bwd")
        off, _ = g1_param_layout(dx, dd)
        views = tuple(grads[o:o + _numel(sh)].view(sh) for o, sh in zip(off, shapes))
        return (None, None, None, None, None, None, None) + views


def _numel(shape):
    n = 1
    for s in shape:
        n *= s
    return n


def g1_mlp(ex, ed, samples_per_ray, dx, dd, precision, flat, params):
    return _G1MLP.apply(ex, ed, samples_per_ray, dx, dd, precision, flat, *params)


# ---------------------------------------------------------------- optimiser ----
def adam_step(flat_param, flat_grad, exp_avg, exp_avg_sq, step, lr=1e-4, betas=(0.9, 0.999), eps=1e-8,
              grad_scale=1.0):
    check(lib().nerf_adam_step(ptr(flat_param), ptr(flat_grad), ptr(exp_avg), ptr(exp_avg_sq), flat_param.numel(),
                               step, lr, betas[0], betas[1], eps, grad_scale, stream()), "nerf_adam_step")


def mse_loss(pixels, target, want_grad=True):
    pixels, target = _f32c(pixels), _f32c(target)
    loss = torch.empty(1, device=pixels.device)
    g = torch.empty_like(pixels) if want_grad else None
    check(lib().nerf_mse_loss(ptr(pixels), ptr(target), pixels.numel(), ptr(loss), ptr(g), stream()), "nerf_mse_loss")
    return loss, g


# ------------------------------------------------------------ tcgen05 path ------
def g1_pack_weights(flat, dx=3, dd=3):
    """fp32 flat parameters -> bf16 UMMA-ready weight image (+ folded fp32 epilogue vectors)."""
    n = lib().nerf_g1_packed_bytes(dx, dd)
    if n <= 0:
        raise ValueError("bad Dx/Dd")
    packed = torch.empty(n, dtype=torch.uint8, device=flat.device)
    check(lib().nerf_g1_pack_weights(ptr(flat), dx, dd, ptr(packed), stream()), "nerf_g1_pack_weights")
    return packed


class _G1TrainTC(torch.autograd.Function):
    """t [B,S], sigma [B,S], rgb [B,S,3] = sample -> encode -> MLP on tcgen05 with every unit's output kept
    as bf16 slab images (`nerf_g1_train_fwd_tc`); backward = the tcgen05 data-gradient kernel, the
    weight-gradient GEMM and the unfold (`nerf_g1_train_bwd_tc`).  This is the node behind
    `pixels = nerf.render_rays(...); loss.backward()` of the reference loop
    (train_conditional_nerf.py:120-135, example.ipynb cell 8) when precision = "bf16": the gradients
    come back as 44 views of one flat buffer in state_dict order."""

    @staticmethod
    def forward(ctx, flat, rays_o, rays_d, t_base, t_in, u, states_x, cfg, *params):
        dx, s, near, far, norm, norm_dir, jitter, seed, ray_base, x3 = cfg
        L = lib()
        b = rays_o.shape[0]
        dev = rays_o.device
        st = stream()
        t = torch.empty(b, s, device=dev); sigma = torch.empty(b, s, device=dev); rgb = torch.empty(b, s, 3, device=dev)
        if x3:                                                  # fp32-grade operands (hi/lo split), unconditional model
            n = L.nerf_g1_train_x3_workspace_bytes(b, s)
            ws = torch.empty(n, dtype=torch.uint8, device=dev)
            packed = g1_pack_weights_x3(flat)
            packed_bwd = torch.empty(L.nerf_g1_packed_bwd_x3_bytes(), dtype=torch.uint8, device=dev)
            check(L.nerf_g1_pack_weights_bwd_x3(ptr(flat), ptr(packed_bwd), st), "nerf_g1_pack_weights_bwd_x3")
            check(L.nerf_g1_train_fwd_x3(ptr(packed), ptr(rays_o), ptr(rays_d), b, s, float(near), float(far), ptr(t_base), ptr(t_in),
                                         ptr(u), int(jitter), seed, ray_base, float(norm), int(norm_dir),
                                         ptr(t), ptr(sigma), ptr(rgb), ptr(ws), n, st), "nerf_g1_train_fwd_x3")
            COUNTERS["g1_train_fwd_x3"] += 1
        else:
            n = L.nerf_g1_train_workspace_bytes(b, s, dx)
            if n <= 0:
                raise ValueError("nerf_g1_train_workspace_bytes: bad argument")
            ws = torch.empty(n, dtype=torch.uint8, device=dev)
            packed = torch.empty(L.nerf_g1_packed_bytes(dx, 3), dtype=torch.uint8, device=dev)
            packed_bwd = torch.empty(L.nerf_g1_packed_bwd_bytes(dx, 3), dtype=torch.uint8, device=dev)
            check(L.nerf_g1_pack_weights(ptr(flat), dx, 3, ptr(packed), st), "nerf_g1_pack_weights")
            check(L.nerf_g1_pack_weights_bwd(ptr(flat), dx, 3, ptr(packed_bwd), st), "nerf_g1_pack_weights_bwd")
            check(L.nerf_g1_train_fwd_tc(ptr(packed), dx, 3, ptr(rays_o), ptr(rays_d), b, s, float(near), float(far),
                                         ptr(t_base), ptr(t_in), ptr(u), int(jitter), seed, ray_base, float(norm), int(norm_dir),
                                         ptr(states_x), ptr(t), ptr(sigma), ptr(rgb), ptr(ws), n, st), "nerf_g1_train_fwd_tc")
            COUNTERS["g1_train_fwd_tc"] += 1
        ctx.save_for_backward(flat, ws, packed, packed_bwd)
        ctx.cfg = (dx, b, s, n, [tuple(q.shape) for q in params], x3)
        ctx.mark_non_differentiable(t)
        return t, sigma, rgb

    @staticmethod
    def backward(ctx, _g_t, d_sigma, d_rgb):
        flat, ws, packed, packed_bwd = ctx.saved_tensors
        dx, b, s, n, shapes, x3 = ctx.cfg
        dev = flat.device
        d_sigma = _f32c(d_sigma) if d_sigma is not None else torch.zeros(b, s, device=dev)
        d_rgb = _f32c(d_rgb) if d_rgb is not None else torch.zeros(b, s, 3, device=dev)
        grads = torch.empty_like(flat)                          # overwritten by the unfold kernel
        if x3:
            check(lib().nerf_g1_train_bwd_x3(ptr(packed), ptr(packed_bwd), ptr(flat), ptr(d_sigma), ptr(d_rgb), b, s,
                                             ptr(grads), ptr(ws), n, 0, stream()), "nerf_g1_train_bwd_x3")
            COUNTERS["g1_train_bwd_x3"] += 1
        else:
            check(lib().nerf_g1_train_bwd_tc(ptr(packed), ptr(packed_bwd), ptr(flat), dx, 3, ptr(d_sigma), ptr(d_rgb), b, s,
                                             ptr(grads), ptr(ws), n, 0, stream()), "nerf_g1_train_bwd_tc")
            COUNTERS["g1_train_bwd_tc"] += 1
        off, _ = g1_param_layout(dx, 3)
        views = tuple(grads[o:o + _numel(sh)].view(sh) for o, sh in zip(off, shapes))
        return (None,) * 8 + views


def g1_train_tc(flat, params, rays_o, rays_d, num_samples, near, far, normalize_position, normalize_direction=True,
                t_base=None, t_in=None, u=None, states_x=None, jitter=False, seed=0, ray_id_base=0, x3=False):
    """Differentiable (w.r.t. the 44 parameters) tcgen05 forward: -> t, sigma [B,S], rgb [B,S,3].
    x3: fp32-grade operands (hi/lo split) in the forward, the data-gradient chain and the weight-gradient GEMMs."""
    states_x = _f32c(states_x)
    dx = 3 if states_x is None else 3 + states_x.shape[-1]
    if x3 and dx != 3:
        raise ValueError("the bf16x3 training path has no conditional inputs")
    cfg = (dx, int(num_samples), near, far, normalize_position, normalize_direction, jitter, seed, ray_id_base, bool(x3))
    return _G1TrainTC.apply(flat, _f32c(rays_o), _f32c(rays_d), _f32c(t_base), _f32c(t_in), _f32c(u), states_x, cfg, *params)


def g1_pack_weights_x3(flat):
    """fp32 flat parameters -> hi/lo bf16 weight image of the fp32-grade tensor-core forward (csrc/fused_tc_x3.cu)."""
    packed = torch.empty(lib().nerf_g1_packed_x3_bytes(), dtype=torch.uint8, device=flat.device)
    check(lib().nerf_g1_pack_weights_x3(ptr(flat), ptr(packed), stream()), "nerf_g1_pack_weights_x3")
    return packed


def g1_mlp_tc_debug(packed, rays_o, rays_d, t, normalize_position, normalize_direction=True, dbg_layer=-1):
    rays_o, rays_d, t = _f32c(rays_o), _f32c(rays_d), _f32c(t)
    b, s = t.shape
    dev = t.device
    sigma = torch.empty(b, s, device=dev); rgb = torch.empty(b, s, 3, device=dev)
    scratch = torch.empty(b, s, device=dev)
    dbg = torch.zeros(b * s, 256, device=dev) if dbg_layer >= 0 else None
    check(lib().nerf_g1_mlp_tc_debug(ptr(packed), ptr(rays_o), ptr(rays_d), ptr(t), b, s, float(normalize_position),
                                     int(normalize_direction), ptr(sigma), ptr(rgb), ptr(scratch), ptr(dbg), dbg_layer,
                                     stream()), "nerf_g1_mlp_tc_debug")
    return sigma, rgb, dbg


def g1_render_fused(packed, num_samples, near, far, normalize_position, normalize_direction=True,
                    rays_o=None, rays_d=None, camera=None, ray_begin=0, num_rays=None,
                    t_base=None, t_in=None, u=None, jitter=False, seed=0, ray_id_base=0, noise_std=0.0,
                    want=("pixels", "depth", "acc", "weights", "t"), states_x=None, rays_per_state=1, x3=False):
    """One fused pass.  camera = (cam_o [n,3], cam_r [n,3,3], H, W, focal) when rays are not given.
    states_x [n,Sx] (conditional model, `packed` built with density_inputs = 3 + Sx): one row per
    `rays_per_state` consecutive rays (camera mode: flattened ray ids)."""
    if rays_o is not None:
        rays_o, rays_d = _f32c(rays_o), _f32c(rays_d)
        b = rays_o.shape[0]; dev = rays_o.device
        cam_o = cam_r = None; n_cams = h = w = 0; focal = 1.0
    else:
        cam_o, cam_r, h, w, focal = camera
        cam_o, cam_r = _f32c(cam_o), _f32c(cam_r)
        n_cams = cam_o.shape[0]; b = num_rays; dev = cam_o.device
    s = num_samples
    out = {}
    alloc = {"pixels": (b, 3), "depth": (b,), "acc": (b,), "weights": (b, s), "t": (b, s), "sigma": (b, s), "rgb": (b, s, 3)}
    for k in want:
        out[k] = torch.empty(*alloc[k], device=dev, dtype=torch.float32)
    states_x = _f32c(states_x)
    n_states, dx = (0, 3) if states_x is None else (states_x.shape[0], 3 + states_x.shape[1])
    nbytes = lib().nerf_g1_render_workspace_bytes(b, s, dx, n_states)
    ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev)
    g = lambda k: ptr(out.get(k))
    if x3:                                                          # fp32-grade operands: `packed` from g1_pack_weights_x3
        if states_x is not None:
            raise ValueError("the bf16x3 forward has no conditional inputs")
        check(lib().nerf_g1_render_fused_x3(ptr(packed), ptr(rays_o), ptr(rays_d), ptr(cam_o), ptr(cam_r), n_cams, h, w,
                                            float(focal), ray_begin, b, s, float(near), float(far), ptr(t_base), ptr(_f32c(t_in)),
                                            ptr(_f32c(u)), int(jitter), seed, ray_id_base, float(noise_std),
                                            float(normalize_position), int(normalize_direction),
                                            g("pixels"), g("depth"), g("acc"), g("weights"), g("t"), g("sigma"), g("rgb"),
                                            ptr(ws), nbytes, stream()), "nerf_g1_render_fused_x3")
        COUNTERS["g1_render_fused_x3"] += 1
        return out
    check(lib().nerf_g1_render_fused(ptr(packed), dx, 3, ptr(rays_o), ptr(rays_d), ptr(cam_o), ptr(cam_r), n_cams, h, w,
                                     float(focal), ray_begin, b, s, float(near), float(far), ptr(t_base), ptr(_f32c(t_in)),
                                     ptr(_f32c(u)), int(jitter), seed, ray_id_base, None, float(noise_std),
                                     float(normalize_position), int(normalize_direction),
                                     ptr(states_x), n_states, int(rays_per_state),
                                     g("pixels"), g("depth"), g("acc"), g("weights"), g("t"), g("sigma"), g("rgb"),
                                     ptr(ws), nbytes, stream()), "nerf_g1_render_fused")
    COUNTERS["g1_render_fused"] += 1
    return out
