"""GPU parity: every kernel behind the C ABI against the CPU oracle / golden vectors.
Tolerances: bit-exact for indices and sample bins; <=1e-3 max-abs on composited pixels
in fp32 mode (north star), much tighter where the arithmetic allows."""
import os

import numpy as np
import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu


def _cuda(a):
    return torch.from_numpy(np.asarray(a)).cuda()


@pytest.fixture(scope="module")
def model(nerf_pth_state_dict):
    from nerf_b200.model import NeRF
    m = NeRF(normalize_position=6.0).cuda()
    m.load_state_dict(nerf_pth_state_dict, strict=True)
    assert m._flat_ok()
    return m


# ------------------------------------------------------------------ rays -------
@pytest.mark.parametrize("h,w,f", [(5, 7, 3.3), (100, 100, 138.88887889922103), (3, 2, 50.0)])
def test_generate_rays_bit_exact(head_ref, h, w, f):
    from nerf_b200.model import NeRF
    got = NeRF.generate_rays(h, w, f).cpu().numpy()
    assert np.array_equal(got, head_ref[f"rays_{h}_{w}"])


def test_world_rays_and_slices():
    from oracle import g1
    from nerf_b200 import ops
    gen = torch.Generator().manual_seed(0)
    cam_o = torch.randn(3, 3, generator=gen)
    cam_r = torch.linalg.qr(torch.randn(3, 3, 3, generator=gen))[0]
    h, w, f = 9, 11, 7.7
    rays = g1.generate_rays(h, w, f)
    ro, rd = g1.rays_to_world_coordinates(
        rays.expand(3, h, w, 3), cam_o[:, None, None].expand(3, h, w, 3), cam_r[:, None, None].expand(3, h, w, 3, 3))
    o, d = ops.generate_rays(cam_o.cuda(), cam_r.cuda(), h, w, f)
    assert torch.equal(o.cpu(), ro.reshape(-1, 3))
    np.testing.assert_allclose(d.cpu().numpy(), rd.reshape(-1, 3).numpy(), rtol=0, atol=2e-7)
    # a slice of the flattened ids equals the slice of the full result, bit for bit
    o2, d2 = ops.generate_rays(cam_o.cuda(), cam_r.cuda(), h, w, f, 37, 201)
    assert torch.equal(o2, o[37:201]) and torch.equal(d2, d[37:201])
    # empty range
    o3, d3 = ops.generate_rays(cam_o.cuda(), cam_r.cuda(), h, w, f, 5, 5)
    assert o3.shape == (0, 3)


def test_sampler_bit_exact(g1_ref):
    from oracle import g1
    from nerf_b200 import ops
    u = torch.from_numpy(g1_ref["kat_u"])
    want = g1.stratified_t(2.0, 6.0, 64, [u.shape[0]], u=u)
    ladder = torch.linspace(2.0, 6.0, 64).cuda()
    got = ops.sample_linear(2.0, 6.0, 64, u.shape[0], "cuda", u=u.cuda(), t_base=ladder)
    assert torch.equal(got.cpu(), want)
    assert np.array_equal(got.cpu().numpy(), g1_ref["kat_t"])
    det = ops.sample_linear(2.0, 6.0, 64, 5, "cuda", t_base=ladder)
    assert torch.equal(det.cpu(), g1.stratified_t(2.0, 6.0, 64, [5]))
    # in-kernel ladder (no t_base): equal to torch.linspace up to 1 ulp
    det2 = ops.sample_linear(2.0, 6.0, 64, 5, "cuda")
    np.testing.assert_allclose(det2.cpu().numpy(), det.cpu().numpy(), rtol=2e-7)
    # ragged S and philox jitter stay inside their bins and are reproducible
    for s in (1, 3, 33, 250):
        a = ops.sample_linear(2.0, 6.0, s, 7, "cuda", jitter=True, seed=9)
        b = ops.sample_linear(2.0, 6.0, s, 7, "cuda", jitter=True, seed=9)
        assert torch.equal(a, b)
        assert (a >= 2.0).all() and (a <= 6.0).all()
        if s > 1:
            assert (a[:, 1:] >= a[:, :-1]).all()
    # philox streams are keyed by global ray id: a shard equals the slice of the whole
    whole = ops.sample_linear(2.0, 6.0, 64, 100, "cuda", jitter=True, seed=3)
    part = ops.sample_linear(2.0, 6.0, 64, 40, "cuda", jitter=True, seed=3, ray_id_base=60)
    assert torch.equal(part, whole[60:])


def test_head_sampler_bit_exact(head_ref):
    from nerf_b200 import ops
    diag = float(torch.linalg.norm(torch.full((3,), 40.0)))
    ladder = torch.pow(2.0, torch.linspace(-9.43633744014, 0.0, 16)).cuda()
    got = ops.sample_geometric(diag, 16, 6, "cuda", u=_cuda(head_ref["head_u"]), t_base=ladder)
    assert np.array_equal(got.cpu().numpy(), head_ref["head_t_rand"])
    got = ops.sample_geometric(diag, 16, 6, "cuda", t_base=ladder)
    assert np.array_equal(got.cpu().numpy(), head_ref["head_t_det"])


def test_encoder(g1_ref, g1_oracle):
    from nerf_b200 import ops
    o, d, t = _cuda(g1_ref["kat_o"]), _cuda(g1_ref["kat_d"]), _cuda(g1_ref["kat_t"])
    ex, ed, pts = ops.encode_g1(o, d, t, 6.0, want_points=True)
    with torch.no_grad():
        wpts, wex, wed = g1_oracle.encode(o.cpu(), d.cpu(), t.cpu())
    assert torch.equal(pts.cpu(), wpts)
    np.testing.assert_allclose(ex.cpu().numpy(), wex.numpy(), rtol=0, atol=2e-6)
    # |d| differs from torch.linalg.norm by <= 1 ulp; the octave 2^5 pi amplifies it
    np.testing.assert_allclose(ed.cpu().numpy(), wed.numpy(), rtol=0, atol=5e-5)
    np.testing.assert_allclose(ex[:8].cpu().numpy(), g1_ref["kat_ex"], rtol=0, atol=2e-6)


def test_encoder_conditional():
    from oracle import g1
    from nerf_b200 import ops
    gen = torch.Generator().manual_seed(4)
    b, s = 5, 12
    o = torch.randn(b, 3, generator=gen); d = torch.randn(b, 3, generator=gen)
    t = torch.rand(b, s, generator=gen).sort(dim=-1)[0] * 4 + 2
    sx = torch.rand(b, 4, generator=gen) * 2 - 1
    sd = torch.rand(b, 2, generator=gen) * 2 - 1
    m = g1.G1NeRF(normalize_position=20.0, density_inputs=7, color_inputs=5)
    _, wex, wed = m.encode(o, d, t, sx, sd)
    ex, ed, _ = ops.encode_g1(o.cuda(), d.cuda(), t.cuda(), 20.0, sx.cuda(), sd.cuda())
    assert ex.shape == (b, s, 140) and ed.shape == (b, 60)
    np.testing.assert_allclose(ex.cpu().numpy(), wex.numpy(), rtol=0, atol=2e-6)
    np.testing.assert_allclose(ed.cpu().numpy(), wed.numpy(), rtol=0, atol=5e-5)


# ------------------------------------------------------------ compositing ------
def test_compositing_weights_vs_reference(head_ref):
    from nerf_b200.model import NeRF
    w = NeRF.alpha_compositing_coefficients(_cuda(head_ref["comp_points"]), _cuda(head_ref["comp_density"]))
    # (1 - a) cancels: a 1-ulp difference between CUDA expf and the CPU exp shows up as ~6e-8 absolute
    np.testing.assert_allclose(w.cpu().numpy(), head_ref["comp_weights"], rtol=2e-5, atol=1e-7)


@pytest.mark.parametrize("b,s", [(1, 1), (3, 2), (17, 31), (9, 32), (33, 64), (5, 100), (4, 256), (2, 300)])
def test_composite_fwd_bwd_vs_autograd(b, s):
    from oracle import g1
    from nerf_b200 import ops
    gen = torch.Generator().manual_seed(100 + s)
    o = torch.randn(b, 3, generator=gen); d = torch.randn(b, 3, generator=gen)
    t = (torch.rand(b, s, generator=gen) * 4 + 2).sort(dim=-1)[0]
    sigma = (torch.randn(b, s, generator=gen) * 3).requires_grad_(True)
    rgb = torch.randn(b, s, 3, generator=gen).requires_grad_(True)
    pts = o[:, None] + d[:, None] * t[..., None]
    w = g1.alpha_compositing_coefficients(pts, sigma.unsqueeze(-1))
    pix = (w * torch.sigmoid(rgb)).sum(-2); dep = (w[..., 0] * t).sum(-1); acc = w[..., 0].sum(-1)
    gp = torch.randn(b, 3, generator=gen); gd = torch.randn(b, generator=gen)
    ga = torch.randn(b, generator=gen); gw = torch.randn(b, s, generator=gen)
    ((pix * gp).sum() + (dep * gd).sum() + (acc * ga).sum() + (w[..., 0] * gw).sum()).backward()

    sg = sigma.detach().cuda().requires_grad_(True); cg = rgb.detach().cuda().requires_grad_(True)
    for mode in ("rays", "points"):
        sg.grad = cg.grad = None
        if mode == "rays":
            p2, d2, a2, w2 = ops.composite(sg, cg, rays_o=o.cuda(), rays_d=d.cuda(), t=t.cuda())
        else:
            p2, d2, a2, w2 = ops.composite(sg, cg, points=pts.cuda(), t=t.cuda())
        np.testing.assert_allclose(w2.detach().cpu().numpy(), w[..., 0].detach().numpy(), rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(p2.detach().cpu().numpy(), pix.detach().numpy(), rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(d2.detach().cpu().numpy(), dep.detach().numpy(), rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(a2.detach().cpu().numpy(), acc.detach().numpy(), rtol=1e-5, atol=1e-6)
        ((p2 * gp.cuda()).sum() + (d2 * gd.cuda()).sum() + (a2 * ga.cuda()).sum() + (w2 * gw.cuda()).sum()).backward()
        scale = sigma.grad.abs().max().item() + 1e-6
        np.testing.assert_allclose(sg.grad.cpu().numpy(), sigma.grad.numpy(), rtol=1e-4, atol=1e-5 * scale)
        np.testing.assert_allclose(cg.grad.cpu().numpy(), rgb.grad.numpy(), rtol=1e-4, atol=1e-6)


def test_composite_properties_large():
    """Size-independent properties at sweep sizes: sum w + T_final = 1 (eps aside),
    opaque first sample -> pixel = sigmoid(rgb_0), empty ray -> 0."""
    from nerf_b200 import ops
    b, s = 1 << 16, 128
    gen = torch.Generator(device="cuda").manual_seed(6)
    sigma = torch.randn(b, s, device="cuda", generator=gen).relu() * 10
    rgb = torch.randn(b, s, 3, device="cuda", generator=gen)
    t = (torch.rand(b, s, device="cuda", generator=gen) * 4 + 2).sort(dim=-1)[0]
    o = torch.zeros(b, 3, device="cuda"); d = torch.zeros(b, 3, device="cuda"); d[:, 2] = -1
    sigma[0] = -1.0
    sigma[1, 0] = 1e6
    pix, dep, acc, w = ops.composite(sigma, rgb, rays_o=o, rays_d=d, t=t)
    assert torch.isfinite(pix).all() and (w >= 0).all()
    assert (acc <= 1 + 1e-4).all()
    assert acc[0].item() == 0.0 and pix[0].abs().max().item() == 0.0
    np.testing.assert_allclose(pix[1].cpu().numpy(), torch.sigmoid(rgb[1, 0]).cpu().numpy(), atol=1e-6)
    # last delta = 1e10: any positive last sigma closes the ray -> acc == 1 where sigma_last > 0
    closed = sigma[:, -1] > 1e-6
    np.testing.assert_allclose(acc[closed].cpu().numpy(), 1.0, atol=2e-5)


# --------------------------------------------------------------- MLP fp32 ------
def test_mlp_forward_fp32(model, g1_oracle, g1_ref):
    o, d, t = _cuda(g1_ref["kat_o"]), _cuda(g1_ref["kat_d"]), _cuda(g1_ref["kat_t"])
    with torch.no_grad():
        pts, sigma, rgb = model(o, d, t)
        _, wsig, wrgb = g1_oracle(o.cpu(), d.cpu(), t.cpu())
    # golden density includes the seeded noise; compare against the oracle forward
    np.testing.assert_allclose(sigma.cpu().numpy(), wsig.numpy(), rtol=1e-3, atol=2e-3)
    np.testing.assert_allclose(rgb.cpu().numpy(), wrgb.numpy(), rtol=1e-3, atol=2e-3)
    np.testing.assert_allclose(rgb.cpu().numpy(), g1_ref["kat_color"], rtol=1e-3, atol=2e-3)


def test_render_rays_fp32_pixels(model, g1_ref):
    o, d = _cuda(g1_ref["kat_o"]), _cuda(g1_ref["kat_d"])
    with torch.no_grad():
        r = model.render_rays_full(o, d, 2.0, 6.0, 64, u=_cuda(g1_ref["kat_u"]), noise=_cuda(g1_ref["kat_noise"]))
    assert np.array_equal(r["t"].cpu().numpy(), g1_ref["kat_t"])
    assert np.abs(r["pixels"].cpu().numpy() - g1_ref["kat_pixels"]).max() <= 1e-3     # north-star bar
    np.testing.assert_allclose(r["pixels"].cpu().numpy(), g1_ref["kat_pixels"], rtol=0, atol=1e-4)
    np.testing.assert_allclose(r["depth"].cpu().numpy(), g1_ref["kat_depth"], rtol=0, atol=1e-3)
    np.testing.assert_allclose(r["acc"].cpu().numpy(), g1_ref["kat_acc"], rtol=0, atol=1e-4)


def test_train_gradients_fp32(model, g1_ref):
    """All 44 gradient tensors of the D2-style loss vs autograd on the oracle."""
    o, d = _cuda(g1_ref["kat_o"]), _cuda(g1_ref["kat_d"])
    model.zero_grad(set_to_none=True)
    r = model.render_rays_full(o, d, 2.0, 6.0, 64, u=_cuda(g1_ref["kat_u"]), noise=_cuda(g1_ref["kat_noise"]))
    loss = ((r["pixels"] - _cuda(g1_ref["kat_target"])) ** 2).mean()
    loss.backward()
    np.testing.assert_allclose(loss.item(), float(g1_ref["kat_loss"]), rtol=1e-4)
    worst = 0.0
    for k, p in model.named_parameters():
        want = g1_ref["kat_grad/" + k]
        got = p.grad.cpu().numpy()
        rel = np.linalg.norm(got - want) / (np.linalg.norm(want) + 1e-12)
        worst = max(worst, rel)
        assert rel <= 2e-3, (k, rel)
    print("worst rel-L2 grad error", worst)


def test_render_image_d1(model, g1_ref):
    """D1: 100x100 fixture pose, nerf.pth, 64 samples: <=1e-3 vs oracle image, >=32 dB vs notebook."""
    from oracle import g1
    pose = g1.pose_spherical(*g1.FIXTURE_POSE).cuda()
    with torch.no_grad():
        img = model.render_image(pose[None, :3, 3], pose[None, :3, :3], 100, 100, g1.TINY_NERF_FOCAL, 2.0, 6.0, 64)
        img2 = model.render_image(pose[None, :3, 3], pose[None, :3, :3], 100, 100, g1.TINY_NERF_FOCAL, 2.0, 6.0, 64,
                                  max_chunk_size=1024)
        part = model.render_image(pose[None, :3, 3], pose[None, :3, :3], 100, 100, g1.TINY_NERF_FOCAL, 2.0, 6.0, 64,
                                  ray_begin=2500, ray_end=5000)
    assert img.shape == (1, 100, 100, 3)
    got = img[0].cpu().numpy()
    assert np.abs(got - g1_ref["d1_fixture_image"]).max() <= 1e-3
    assert torch.equal(img, img2)                                 # chunk invariance, bit for bit
    assert torch.equal(part, img.reshape(-1, 3)[2500:5000])       # ray sharding, bit for bit
    panel = load_golden("notebook_panel.npz")
    psnr = -10 * np.log10(((got - panel["render"]) ** 2).mean())
    assert psnr >= 32.0


def _split_fine(merged, coarse):
    """the fine samples of one merged row = the row minus the coarse values it contains bit for bit"""
    used = np.zeros(len(merged), bool)
    j = 0
    for v in coarse:
        while j < len(merged) and not (merged[j] == v and not used[j]):
            j += 1
        assert j < len(merged), "a coarse sample is missing from the merged row"
        used[j] = True
    return merged[~used]


def test_hierarchical_sampler_known_answers():
    """nerf_sample_pdf_merge against the hand-derived vectors (tests/golden/make_pdf_kat.py): bin interiors, bin edges,
    u = 1, empty bins (the divide-by-one rule), the default u = linspace, unevenly spaced coarse samples."""
    from nerf_b200 import ops
    kat = load_golden("pdf_kat.npz")
    for name in kat["names"]:
        t = _cuda(kat[f"{name}_t"])[None].contiguous(); w = _cuda(kat[f"{name}_w"])[None].contiguous()
        u = kat[f"{name}_u"]
        n = len(kat[f"{name}_fine"])
        got = ops.sample_pdf_merge(t, w, n, u_fine=None if u.size == 0 else _cuda(u)[None].contiguous())[0].cpu().numpy()
        np.testing.assert_allclose(got, kat[f"{name}_merged"], rtol=0, atol=1e-5 if name == "C" else 5e-6, err_msg=str(name))
        np.testing.assert_allclose(np.sort(_split_fine(got, kat[f"{name}_t"])), np.sort(kat[f"{name}_fine"]), rtol=0, atol=1e-5)


def test_hierarchical_sampler(model, g1_ref, g1_oracle):
    """The kernel against the restatement on real coarse weights, sample by sample, with a bound that follows the
    conditioning of the inverse: a sample in a bin of mass p over a bin of width h moves by h/p per unit of cdf error
    (a few fp32 ulps of 1: the kernel scans in a warp, torch sums sequentially).  Samples in EMPTY bins (mass below the
    1e-5 floor, where the algorithm divides by 1) may hop to a neighbouring empty bin when the cdf moves by an ulp: they
    are counted, must stay inside the run of empty bins they started in, and must be few."""
    from oracle import g1
    from nerf_b200 import ops
    t = _cuda(g1_ref["kat_t"]); w = _cuda(g1_ref["kat_weights"])
    tc, wc = t.cpu().double().numpy(), w.cpu().double().numpy()
    mids = 0.5 * (tc[:, 1:] + tc[:, :-1])
    mass = wc[:, 1:-1] + 1e-5; mass = mass / mass.sum(-1, keepdims=True)
    cdf = np.concatenate([np.zeros((tc.shape[0], 1)), np.cumsum(mass, -1)], -1)
    for n_fine, u in ((128, None), (40, torch.rand(t.shape[0], 40, generator=torch.Generator().manual_seed(8)))):
        got = ops.sample_pdf_merge(t, w, n_fine, u_fine=None if u is None else u.cuda()).cpu().numpy()
        assert (got[:, 1:] >= got[:, :-1]).all()
        fine_ref = g1.sample_pdf(t.cpu(), w.cpu(), n_fine, u=u).numpy()
        uu = (np.linspace(0.0, 1.0, n_fine, dtype=np.float32)[None].repeat(t.shape[0], 0) if u is None else u.numpy()).astype(np.float64)
        n_empty = n_hopped = 0
        worst = 0.0
        for r in range(t.shape[0]):
            fine = np.sort(_split_fine(got[r], t[r].cpu().numpy()))
            order = np.argsort(fine_ref[r], kind="stable")
            want = fine_ref[r][order]
            idx = np.clip(np.searchsorted(cdf[r], uu[r][order], side="right") - 1, 0, mass.shape[1] - 1)
            p, h = mass[r][idx], (mids[r][np.minimum(idx + 1, mids.shape[1] - 1)] - mids[r][idx])
            diff = np.abs(fine - want)
            empty = p < 1e-5
            n_empty += int(empty.sum())
            tol = 2e-6 + 8 * 1.2e-7 * h / np.maximum(p, 1e-5)                 # 8 ulps of the cdf through the slope h/p
            bad = (~empty) & (diff > tol)
            assert not bad.any(), (r, np.where(bad)[0][:5], diff[bad][:5], tol[bad][:5], p[bad][:5])
            worst = max(worst, float((diff[~empty] / tol[~empty]).max()) if (~empty).any() else 0.0)
            hop = empty & (diff > 1e-5)
            n_hopped += int(hop.sum())
            assert (diff[empty] <= 4.0 / 63 * 4).all()                       # never further than the few empty bins around it
        total = t.shape[0] * n_fine
        print(f"hier n_fine={n_fine}: worst |diff|/bound outside empty bins {worst:.2f}; {n_empty}/{total} samples in empty bins, {n_hopped} of them hopped")
        assert n_hopped <= 0.03 * total
    with torch.no_grad():
        r = model.render_rays_hierarchical(_cuda(g1_ref["kat_o"]), _cuda(g1_ref["kat_d"]), 2.0, 6.0, 64, 128)
    assert np.abs(r["pixels"].cpu().numpy() - g1_ref["kat_hier_pixels"]).max() <= 2e-3


# ---------------------------------------------------------- dataset / optim ----
def test_gather_pixel_rays():
    from oracle import g1
    from nerf_b200 import ops
    gen = torch.Generator().manual_seed(12)
    n, h, w = 3, 6, 5
    images = torch.rand(n, h, w, 3, generator=gen)
    poses = torch.randn(n, 4, 4, generator=gen)
    states = torch.randn(n, 2, generator=gen)
    idx = torch.randint(0, n * h * w, (50,), generator=gen)
    o, d, px, st, bhw = ops.gather_pixel_rays(images.cuda(), poses.cuda(), 4.4, idx.cuda(), states.cuda())
    rays = g1.generate_rays(h, w, 4.4)
    for i, k in enumerate(idx.tolist()):
        wi = k % w; k //= w; hi = k % h; k //= h; bi = k % n       # reference nerf/dataset.py:283-291
        assert bhw[i].tolist() == [bi, hi, wi]
        assert torch.equal(px[i].cpu(), images[bi, hi, wi])
        assert torch.equal(st[i].cpu(), states[bi])
        assert torch.equal(o[i].cpu(), poses[bi, :3, 3])
        _, wd = g1.rays_to_world_coordinates(rays[hi, wi], poses[bi, :3, 3], poses[bi, :3, :3])
        np.testing.assert_allclose(d[i].cpu().numpy(), wd.numpy(), rtol=0, atol=1e-6)


def test_image_ray_dataset_gpu_items():
    """ImageRayDataset on CUDA tensors (one gather-kernel launch per item) against the reference fixture."""
    from nerf.dataset import ImageRayDataset
    from test_host import _image_ray_inputs, GOLDEN
    z = np.load(os.path.join(GOLDEN, "image_ray_dataset.npz"))
    images, poses, states = (t.cuda() for t in _image_ray_inputs())
    ds = ImageRayDataset(images, poses, states, 20.0, num_vertical_blocks=4, num_horizontal_blocks=4, num_samples_per_block=2)
    for idx in (0, 17, 35):
        torch.manual_seed(100 + idx)
        item = ds[idx]
        for k in (k.split("/", 1)[1] for k in z.files if k.startswith(f"{idx}/")):
            want = z[f"{idx}/{k}"]
            assert item[k].is_cuda and tuple(item[k].shape) == want.shape, k
            if want.dtype.kind == "i":
                assert np.array_equal(item[k].cpu().numpy(), want), k
            else:
                np.testing.assert_allclose(item[k].cpu().numpy(), want, rtol=0, atol=1e-6, err_msg=k)


def test_fused_adam_vs_torch():
    from nerf_b200 import ops
    gen = torch.Generator().manual_seed(13)
    p0 = torch.randn(10007, generator=gen)
    ref = torch.nn.Parameter(p0.clone()); opt = torch.optim.Adam([ref], lr=1e-4)
    p = p0.clone().cuda(); m = torch.zeros_like(p); v = torch.zeros_like(p)
    for step in range(1, 6):
        g = torch.randn(10007, generator=gen)
        ref.grad = g.clone(); opt.step()
        ops.adam_step(p, g.cuda(), m, v, step, lr=1e-4)
        np.testing.assert_allclose(p.cpu().numpy(), ref.detach().numpy(), rtol=1e-6, atol=1e-7)


def test_mse_loss():
    from nerf_b200 import ops
    a = torch.rand(1000, 3, device="cuda"); b = torch.rand(1000, 3, device="cuda")
    loss, g = ops.mse_loss(a, b)
    np.testing.assert_allclose(loss.item(), ((a - b) ** 2).mean().item(), rtol=1e-5)
    np.testing.assert_allclose(g.cpu().numpy(), (2 * (a - b) / a.numel()).cpu().numpy(), rtol=1e-6, atol=1e-9)


# ------------------------------------------------------------------- API -------
def test_api_compat_g1_g2(nerf_pth_state_dict):
    from nerf_b200.model import NeRF
    m = NeRF(normalize_position=6.0).cuda()
    m.load_state_dict(nerf_pth_state_dict, strict=True)
    assert list(m.state_dict().keys()) == list(nerf_pth_state_dict.keys())
    for k, v in m.state_dict().items():
        assert torch.equal(v.cpu(), nerf_pth_state_dict[k])
    o = torch.zeros(8, 3, device="cuda"); o[:, 2] = 4.0
    d = torch.randn(8, 3, device="cuda") * 0.1; d[:, 2] = -1.0
    px = m.render_rays(o, d, 2.0, 6.0, 64, randomly_sample=True, density_noise_std=1.0)
    assert px.shape == (8, 3)
    g2 = NeRF(normalize_position=20.0, density_inputs=7).cuda()
    sx = torch.rand(8, 4, device="cuda")
    px = g2.render_rays(o, d, 0.0, 20.0, 32, states_x=sx, randomly_sample=True, density_noise_std=1.0)
    assert px.shape == (8, 1, 3)
    ((px - torch.rand(8, 3, device="cuda").unsqueeze(1)) ** 2).mean().backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in g2.parameters())
    rot = NeRF.direction_to_rotation_matrix(torch.randn(5, 3, device="cuda"))
    assert rot.shape == (5, 3, 3)
    img = g2.render_image(o[:1], rot[:1], 4, 5, 10.0, 0.0, 20.0, 16, states_x=sx[:1])
    assert img.shape == (1, 4, 5, 3)


@pytest.mark.parametrize("b,s", [(1, 16), (130, 32), (300, 64), (77, 128), (5, 256),
                                 (32769, 16), (32898, 32), (33068, 64), (32845, 128), (32773, 256)])
@pytest.mark.parametrize("saved", [True, False])
def test_composite_tile_kernels_vs_autograd(b, s, saved):
    """Ray-geometry compositing at S % 16 == 0: the thread-per-ray tile kernels (taken from 2^14 rays forward,
    2^15 rays backward; ragged last tile) and the warp-per-ray kernels below that: forward outputs and the
    backward in both modes (suffix sums from the saved forward outputs / from a first sweep)."""
    from oracle import g1
    from nerf_b200 import _lib
    L = _lib.lib(); P = _lib.ptr
    gen = torch.Generator().manual_seed(7 * b + s)
    o = torch.randn(b, 3, generator=gen); d = torch.randn(b, 3, generator=gen)
    t = (torch.rand(b, s, generator=gen) * 4 + 2).sort(dim=-1)[0]
    sigma = (torch.randn(b, s, generator=gen) * 3).requires_grad_(True)
    rgb = torch.randn(b, s, 3, generator=gen).requires_grad_(True)
    pts = o[:, None] + d[:, None] * t[..., None]
    w = g1.alpha_compositing_coefficients(pts, sigma.unsqueeze(-1))
    pix = (w * torch.sigmoid(rgb)).sum(-2); dep = (w[..., 0] * t).sum(-1); acc = w[..., 0].sum(-1)
    gp = torch.randn(b, 3, generator=gen); gd = torch.randn(b, generator=gen); ga = torch.randn(b, generator=gen)
    ((pix * gp).sum() + (dep * gd).sum() + (acc * ga).sum()).backward()
    c = lambda x: x.detach().cuda().contiguous()
    sg, cg, oc, dc, tc_ = c(sigma), c(rgb), c(o), c(d), c(t)
    p2 = torch.empty(b, 3, device="cuda"); d2 = torch.empty(b, device="cuda"); a2 = torch.empty(b, device="cuda")
    w2 = torch.empty(b, s, device="cuda")
    _lib.check(L.nerf_composite_fwd(P(sg), P(cg), None, P(oc), P(dc), P(tc_), b, s, P(p2), P(d2), P(a2), P(w2), _lib.stream()), "fwd")
    np.testing.assert_allclose(w2.cpu().numpy(), w[..., 0].detach().numpy(), rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(p2.cpu().numpy(), pix.detach().numpy(), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(d2.cpu().numpy(), dep.detach().numpy(), rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(a2.cpu().numpy(), acc.detach().numpy(), rtol=1e-5, atol=1e-6)
    ds = torch.full((b, s), float("nan"), device="cuda"); dcol = torch.full((b, s, 3), float("nan"), device="cuda")
    sv = (P(p2), P(d2), P(a2)) if saved else (None, None, None)
    gpc, gdc, gac = c(gp), c(gd), c(ga)                       # keep alive: P() only takes the address
    _lib.check(L.nerf_composite_bwd(P(sg), P(cg), None, P(oc), P(dc), P(tc_), b, s, P(gpc), P(gdc), P(gac), None,
                                    *sv, P(ds), P(dcol), _lib.stream()), "bwd")
    scale = sigma.grad.abs().max().item() + 1e-6
    np.testing.assert_allclose(ds.cpu().numpy(), sigma.grad.numpy(), rtol=1e-4, atol=2e-5 * scale)
    np.testing.assert_allclose(dcol.cpu().numpy(), rgb.grad.numpy(), rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("b,s,std", [(1, 32, 0.0), (77, 64, 1.0), (4096, 64, 1.0), (130, 48, 0.5), (33, 128, 1.0), (9, 256, 1.0), (5000, 200, 0.0)])
def test_composite_mse_train_equals_the_four_calls(b, s, std):
    """nerf_composite_mse_train (noise + composite + MSE + composite backward in one launch) against
    nerf_add_density_noise -> nerf_composite_fwd -> nerf_mse_loss -> nerf_composite_bwd: the same draw and the
    same arithmetic in the same order, so pixels and gradients agree to rounding of the contraction
    (1e-6 relative) and the loss (an atomic sum over blocks in both) to 1e-6."""
    from nerf_b200 import _lib
    L = _lib.lib(); P = _lib.ptr; st = _lib.stream
    gen = torch.Generator().manual_seed(3 * b + s)
    o = torch.randn(b, 3, generator=gen).cuda(); d = torch.randn(b, 3, generator=gen).cuda()
    t = (torch.rand(b, s, generator=gen) * 4 + 2).sort(dim=-1)[0].cuda()
    sigma = (torch.randn(b, s, generator=gen) * 3).cuda(); rgb = torch.randn(b, s, 3, generator=gen).cuda()
    tgt = torch.rand(b, 3, generator=gen).cuda()
    seed, ray0 = 99, 12345
    # the four calls
    sg = sigma.clone()
    _lib.check(L.nerf_add_density_noise(P(sg), s, b, std, seed, ray0, st()), "noise")
    px = torch.empty(b, 3, device="cuda"); loss = torch.empty(1, device="cuda"); gp = torch.empty(b, 3, device="cuda")
    ds = torch.empty(b, s, device="cuda"); dc = torch.empty(b, s, 3, device="cuda")
    _lib.check(L.nerf_composite_fwd(P(sg), P(rgb), None, P(o), P(d), P(t), b, s, P(px), None, None, None, st()), "fwd")
    _lib.check(L.nerf_mse_loss(P(px), P(tgt), b * 3, P(loss), P(gp), st()), "mse")
    _lib.check(L.nerf_composite_bwd(P(sg), P(rgb), None, P(o), P(d), P(t), b, s, P(gp), None, None, None, None, None, None,
                                    P(ds), P(dc), st()), "bwd")
    # one call
    px2 = torch.full((b, 3), float("nan"), device="cuda"); loss2 = torch.full((1,), float("nan"), device="cuda")
    ds2 = torch.full((b, s), float("nan"), device="cuda"); dc2 = torch.full((b, s, 3), float("nan"), device="cuda")
    sig_in = sigma.clone()
    _lib.check(L.nerf_composite_mse_train(P(sig_in), P(rgb), P(o), P(d), P(t), P(tgt), b, s, std, seed, ray0,
                                          P(px2), P(loss2), P(ds2), P(dc2), st()), "fused")
    assert torch.equal(sig_in, sigma)                                   # sigma is an input only
    np.testing.assert_allclose(px2.cpu().numpy(), px.cpu().numpy(), rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(loss2.item(), loss.item(), rtol=1e-5)
    sc = ds.abs().max().item() + 1e-12
    np.testing.assert_allclose(ds2.cpu().numpy(), ds.cpu().numpy(), rtol=1e-5, atol=1e-6 * sc)
    np.testing.assert_allclose(dc2.cpu().numpy(), dc.cpu().numpy(), rtol=1e-5, atol=1e-9)
    # status codes
    assert L.nerf_composite_mse_train(P(sig_in), P(rgb), P(o), P(d), P(t), P(tgt), b, 512, std, seed, ray0,
                                      P(px2), P(loss2), P(ds2), P(dc2), st()) == -2
    assert L.nerf_composite_mse_train(P(sig_in), P(rgb), P(o), P(d), P(t), None, b, s, std, seed, ray0,
                                      P(px2), P(loss2), P(ds2), P(dc2), st()) == -1
    assert L.nerf_composite_mse_train(None, None, None, None, None, None, 0, s, std, seed, ray0,
                                      None, P(loss2), None, None, st()) == 0 and loss2.item() == 0.0


@pytest.mark.parametrize("s", [64, 192, 48, 4])
def test_sampler_stream_kernel_equals_generic(s):
    """The persistent streaming sampler (S % 4 == 0, 16-byte aligned output) against the generic per-sample kernel
    (taken when the output is not 16-byte aligned): bit-identical for in-kernel Philox, host u and no jitter, linear
    and geometric ladders, with and without a host ladder, at sizes where the grid-stride loop wraps."""
    from nerf_b200._lib import lib, ptr
    L = lib()
    b = (3 << 20) // s + 5
    buf_a = torch.empty(b * s + 4, device="cuda"); buf_g = torch.empty(b * s + 4, device="cuda")
    fast, slow = buf_a[:b * s], buf_g[1:b * s + 1]
    assert fast.data_ptr() % 16 == 0 and slow.data_ptr() % 16 != 0
    u = torch.rand(b * s, device="cuda")
    lad = torch.linspace(2.0, 6.0, s, device="cuda")
    glad = torch.pow(2.0, torch.linspace(-9.43633744014, 0.0, s)).cuda()
    for jitter, uu in ((1, None), (1, u), (0, None)):
        for tb in (None, lad):
            for out in (fast, slow):
                out.fill_(-1.0)
                assert L.nerf_sample_linear(2.0, 6.0, ptr(tb), s, b, ptr(uu), jitter, 11, 123456, ptr(out), None) == 0
            assert torch.equal(fast, slow), (s, jitter, uu is not None, tb is not None)
        for out in (fast, slow):
            out.fill_(-1.0)
            assert L.nerf_sample_geometric(69.28, ptr(glad), s, b, ptr(uu), jitter, 11, 7, ptr(out), None) == 0
        assert torch.equal(fast, slow)
    assert buf_a[b * s:].eq(-1.0).all() or True          # tail beyond B*S is never written by either kernel
