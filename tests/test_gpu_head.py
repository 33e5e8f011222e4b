"""HEAD generation (the model whose source is in the reference tree) against outputs of the UNMODIFIED
reference stored in tests/golden/head_ref.npz (tests/golden/make_golden.py imports /root/reference)."""
import numpy as np
import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu


def _cuda(a):
    return torch.from_numpy(np.asarray(a)).cuda()


@pytest.fixture(scope="module")
def head_ref():
    return load_golden("head_ref.npz")


@pytest.fixture(scope="module")
def head_model(head_ref):
    from nerf.model import NeRF
    from nerf_b200.head import HeadNeRF
    m = NeRF().cuda()                                           # the reference's default construction
    assert isinstance(m, HeadNeRF) and isinstance(m, NeRF)
    sd = {str(k): torch.from_numpy(head_ref["head_sd/" + str(k)]) for k in head_ref["head_keys"]}
    assert list(m.state_dict().keys()) == list(sd.keys())
    m.load_state_dict(sd, strict=True)
    return m


def test_head_sampler(head_model, head_ref):
    o, d = _cuda(head_ref["head_ro"]), _cuda(head_ref["head_rd"])
    t = head_model.sample_along_rays(o, d, 16, randomly_sample=False)
    assert np.array_equal(t.cpu().numpy(), head_ref["head_t_det"])
    t = head_model.sample_along_rays(o, d, 16, u=_cuda(head_ref["head_u"]))
    assert np.array_equal(t.cpu().numpy(), head_ref["head_t_rand"])


def test_head_integrated_pe(head_model, head_ref):
    """a5-H: conical-frustum Gaussians and the integrated positional encoding."""
    o, d, t = _cuda(head_ref["head_ro"]), _cuda(head_ref["head_rd"]), _cuda(head_ref["head_t_rand"])
    mean, cov, feat = head_model.integrated_pe(o, d, t)
    np.testing.assert_allclose(mean.cpu().numpy(), head_ref["head_ipe_mean"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(cov.cpu().numpy(), head_ref["head_ipe_cov"], rtol=1e-5, atol=1e-9)
    # features with |argument| up to mean * 2^11 ~ 1e5: one ulp of the argument is ~1e-2 rad, so the
    # comparison is only meaningful where the amplitude exp(-var/2) has not already damped the feature
    np.testing.assert_allclose(feat.cpu().numpy(), head_ref["head_ipe_feat"], rtol=0, atol=2e-5)


def test_head_forward(head_model, head_ref):
    """a6-H/a7: trunk outputs (density, colour, segmentation logits)."""
    o, d, t = _cuda(head_ref["head_ro"]), _cuda(head_ref["head_rd"]), _cuda(head_ref["head_t_rand"])
    mean, dens, col, seg = head_model(o, d, t)
    assert dens.shape == (6, 15, 1) and col.shape == (6, 15, 3) and seg.shape == (6, 15, 50)
    np.testing.assert_allclose(mean.cpu().numpy(), head_ref["head_ipe_mean"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(dens.cpu().numpy(), head_ref["head_fwd_density"], rtol=0, atol=1e-4)
    np.testing.assert_allclose(col.cpu().numpy(), head_ref["head_fwd_color"], rtol=0, atol=1e-4)
    np.testing.assert_allclose(seg.cpu().numpy(), head_ref["head_fwd_seg"], rtol=0, atol=1e-4)


def test_head_render_rays(head_model, head_ref):
    """a9-a11: colour composite (<= 1e-3 bar) and the segmentation logsumexp composite."""
    o, d = _cuda(head_ref["head_ro"]), _cuda(head_ref["head_rd"])
    img, seg = head_model.render_rays(o, d, 16, randomly_sample=False, density_noise_std=0.0)
    assert img.shape == (6, 1, 3) and seg.shape == (6, 1, 50)
    assert img.requires_grad            # like the reference: outside no_grad the render carries its autograd graph
    img, seg = img.detach(), seg.detach()
    err = np.abs(img.cpu().numpy() - head_ref["head_rr_img"]).max()
    print("HEAD render_rays pixel max|diff|", err)
    assert err <= 1e-3
    np.testing.assert_allclose(img.cpu().numpy(), head_ref["head_rr_img"], rtol=0, atol=1e-5)
    np.testing.assert_allclose(seg.cpu().numpy(), head_ref["head_rr_seg"], rtol=1e-5, atol=1e-4)


def test_head_render_image(head_model, head_ref):
    """a1-a3 + the whole HEAD chain through render_image (the reference ran it with max_chunk_size=7)."""
    co, cr = _cuda(head_ref["head_ri_co"]), _cuda(head_ref["head_ri_cr"])
    img, seg = head_model.render_image(co, cr, 6, 5, 7.5, 12, max_chunk_size=7)
    assert img.shape == (2, 6, 5, 3) and seg.shape == (2, 6, 5, 50)
    np.testing.assert_allclose(img.cpu().numpy(), head_ref["head_ri_img"], rtol=0, atol=1e-5)
    np.testing.assert_allclose(seg.cpu().numpy(), head_ref["head_ri_seg"], rtol=1e-5, atol=1e-4)
    part_img, part_seg = head_model.render_image(co, cr, 6, 5, 7.5, 12, ray_begin=7, ray_end=41)
    assert torch.equal(part_img, img.reshape(-1, 3)[7:41]) and torch.equal(part_seg, seg.reshape(-1, 50)[7:41])


def test_head_vs_oracle_large(head_model, head_ref):
    """Sizes beyond the fixture (ragged tiles, several segments per ray) against the pinned oracle."""
    from oracle import head as oh
    om = oh.HeadNeRF()
    om.load_state_dict({str(k): torch.from_numpy(head_ref["head_sd/" + str(k)]) for k in head_ref["head_keys"]})
    g = torch.Generator().manual_seed(21)
    o = torch.randn(77, 3, generator=g) * 2; d = torch.randn(77, 3, generator=g)
    u = torch.rand(77, 65, generator=g); noise = torch.randn(77, 64, 1, generator=g)
    with torch.no_grad():
        wi, ws = om.render_rays(o, d, 65, u=u, noise=noise)
    with torch.no_grad():
        gi, gs = head_model.render_rays(o.cuda(), d.cuda(), 65, randomly_sample=True, u=u.cuda(), noise=noise.cuda(), density_noise_std=1.0)
    assert not gi.requires_grad
    assert np.abs(gi.cpu().numpy() - wi.numpy()).max() <= 1e-3
    np.testing.assert_allclose(gi.cpu().numpy(), wi.numpy(), rtol=0, atol=2e-5)
    np.testing.assert_allclose(gs.cpu().numpy(), ws.numpy(), rtol=1e-4, atol=1e-3)


# ------------------------------------------------------------ bf16 tensor-core path ----
@pytest.fixture(scope="module")
def head_model_bf16(head_ref):
    from nerf.model import NeRF
    m = NeRF(encoding_size=32, precision="bf16").cuda()
    m.load_state_dict({str(k): torch.from_numpy(head_ref["head_sd/" + str(k)]) for k in head_ref["head_keys"]}, strict=True)
    return m


def test_head_tc_forward_vs_reference(head_model_bf16, head_ref):
    """tcgen05 encoder + trunk (bf16 operands, fp32 accumulate) against the unmodified reference's outputs:
    Gaussian means in fp32, trunk outputs to bf16 accuracy (activations are O(1) after LayerNorm, six
    chained bf16 GEMMs)."""
    o, d, t = _cuda(head_ref["head_ro"]), _cuda(head_ref["head_rd"]), _cuda(head_ref["head_t_rand"])
    mean, dens, col, seg = head_model_bf16(o, d, t)
    assert dens.shape == (6, 15, 1) and col.shape == (6, 15, 3) and seg.shape == (6, 15, 50)
    np.testing.assert_allclose(mean.cpu().numpy(), head_ref["head_ipe_mean"], rtol=1e-5, atol=1e-5)
    for got, key in ((dens, "head_fwd_density"), (col, "head_fwd_color"), (seg, "head_fwd_seg")):
        want = head_ref[key]
        err = np.abs(got.cpu().numpy() - want).max()
        print(key, "max|diff|", err, "scale", np.abs(want).max())
        assert err <= 3e-2 * max(1.0, np.abs(want).max())


def test_head_tc_render_vs_fp32(head_model, head_model_bf16):
    """Rendered pixels and segmentation of the bf16 path against the fp32 path (itself pinned to the reference at
    1e-5) on ragged sizes: several tiles, rays straddling tiles, a partial last tile."""
    g = torch.Generator().manual_seed(31)
    o = (torch.randn(301, 3, generator=g) * 2).cuda(); d = torch.randn(301, 3, generator=g).cuda()
    u = torch.rand(301, 48, generator=g).cuda()
    with torch.no_grad():
        wi, ws = head_model.render_rays(o, d, 48, randomly_sample=True, u=u)
    gi, gs = head_model_bf16.render_rays(o, d, 48, randomly_sample=True, u=u)
    assert gi.shape == wi.shape and gs.shape == ws.shape
    err = (gi - wi).abs().max().item()
    mse = ((gi - wi) ** 2).mean().item()
    print("HEAD bf16 vs fp32 pixels: max|diff|", err, "PSNR", -10 * np.log10(max(mse, 1e-20)))
    assert err <= 2e-2 and -10 * np.log10(max(mse, 1e-20)) >= 45.0
    assert (gs - ws).abs().max().item() <= 0.15                      # log-probabilities of 50 classes


def test_head_tc_render_image_shards(head_model_bf16):
    """render_image through the bf16 path: shapes, determinism, and a ray shard equals the slice of the full frame."""
    from oracle.g1 import pose_spherical
    pose = pose_spherical(40.0, -30.0, 4.0).cuda()
    co, cr = pose[None, :3, 3].contiguous(), pose[None, :3, :3].contiguous()
    img, seg = head_model_bf16.render_image(co, cr, 20, 24, 30.0, 33)
    assert img.shape == (1, 20, 24, 3) and seg.shape == (1, 20, 24, 50) and torch.isfinite(img).all() and torch.isfinite(seg).all()
    img2, _ = head_model_bf16.render_image(co, cr, 20, 24, 30.0, 33)
    assert torch.equal(img, img2)
    part, _ = head_model_bf16.render_image(co, cr, 20, 24, 30.0, 33, ray_begin=100, ray_end=333)
    assert torch.allclose(part, img.reshape(-1, 3)[100:333], atol=1e-6)


# ------------------------------------------------------------------ fp32 backward ----
def _rel_l2(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def _head_loss(img, seg, target, labels):
    return ((img[:, 0] - target) ** 2).mean() + torch.nn.functional.nll_loss(seg[:, 0], labels)


def test_head_backward_vs_reference(head_model, head_ref):
    """loss.backward() through render_rays (composite + trunk backward kernels) against the parameter gradients
    the UNMODIFIED reference's autograd produced (tests/golden/head_grad_ref.npz): deterministic sampling, and
    stratified sampling + density noise with the reference's own torch.rand / torch.randn draws."""
    gr = load_golden("head_grad_ref.npz")
    o, d = _cuda(head_ref["head_ro"]), _cuda(head_ref["head_rd"])
    target, labels = _cuda(gr["target"]), _cuda(gr["labels"])
    cases = (("det", dict(randomly_sample=False, density_noise_std=0.0)),
             ("noisy", dict(randomly_sample=True, u=_cuda(gr["noisy_u"]), noise=_cuda(gr["noisy_noise"]), density_noise_std=1.0)))
    for tag, kw in cases:
        head_model.zero_grad()
        img, seg = head_model.render_rays(o, d, 16, **kw)
        assert img.requires_grad and seg.requires_grad and img.shape == (6, 1, 3) and seg.shape == (6, 1, 50)
        np.testing.assert_allclose(img.detach().cpu().numpy(), gr[tag + "_img"], rtol=0, atol=1e-5)
        np.testing.assert_allclose(seg.detach().cpu().numpy(), gr[tag + "_seg"], rtol=1e-5, atol=1e-4)
        loss = _head_loss(img, seg, target, labels)
        loss.backward()
        np.testing.assert_allclose(loss.item(), gr[tag + "_loss"], rtol=1e-5)
        worst = 0.0
        for k, p in head_model.named_parameters():
            want = gr[f"{tag}_grad/{k}"]
            assert p.grad is not None and p.grad.shape == want.shape, k
            worst = max(worst, _rel_l2(p.grad.cpu().numpy(), want))
            assert _rel_l2(p.grad.cpu().numpy(), want) <= 2e-3, (tag, k, _rel_l2(p.grad.cpu().numpy(), want))
        print(f"HEAD backward [{tag}] worst per-tensor rel-L2 vs the reference's autograd: {worst:.2e}")
    head_model.zero_grad()


@pytest.mark.parametrize("branch", ["both", "colour", "segmentation"])
def test_head_backward_vs_oracle_large(head_model, head_ref, branch):
    """Ragged sizes (77 rays x 64 intervals: two 32-sample segments per ray, partial row tiles) against autograd
    through the pinned oracle, with either output's gradient absent (NULL d_pixels / d_seg through the ABI)."""
    from oracle import head as oh
    om = oh.HeadNeRF()
    om.load_state_dict({str(k): torch.from_numpy(head_ref["head_sd/" + str(k)]) for k in head_ref["head_keys"]})
    g = torch.Generator().manual_seed(23)
    o = torch.randn(77, 3, generator=g) * 2; d = torch.randn(77, 3, generator=g)
    u = torch.rand(77, 65, generator=g); noise = torch.randn(77, 64, 1, generator=g)
    target = torch.rand(77, 3, generator=g); labels = torch.randint(0, 50, (77,), generator=g)

    def loss_of(img, seg, tgt, lab):
        a = ((img[:, 0] - tgt) ** 2).mean() if branch != "segmentation" else 0.0
        b = torch.nn.functional.nll_loss(seg[:, 0], lab) if branch != "colour" else 0.0
        return a + b

    wi, ws = om.render_rays(o, d, 65, u=u, noise=noise)
    loss_of(wi, ws, target, labels).backward()
    head_model.zero_grad()
    gi, gs = head_model.render_rays(o.cuda(), d.cuda(), 65, randomly_sample=True, u=u.cuda(), noise=noise.cuda(), density_noise_std=1.0)
    loss_of(gi, gs, target.cuda(), labels.cuda()).backward()
    for (k, p), (_, q) in zip(head_model.named_parameters(), om.named_parameters()):
        err = _rel_l2(p.grad.cpu().numpy(), q.grad.numpy())
        assert err <= 2e-3, (branch, k, err)
    head_model.zero_grad()


def test_head_training_reduces_loss(head_ref):
    """A few Adam steps through the public API (model.render_rays -> loss.backward -> torch.optim.Adam.step)."""
    from nerf.model import NeRF
    torch.manual_seed(0)
    m = NeRF().cuda()
    m.load_state_dict({str(k): torch.from_numpy(head_ref["head_sd/" + str(k)]) for k in head_ref["head_keys"]}, strict=True)
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    g = torch.Generator().manual_seed(5)
    o = (torch.randn(512, 3, generator=g) * 2).cuda(); d = torch.randn(512, 3, generator=g).cuda()
    target = torch.rand(512, 3, generator=g).cuda() * 0.2 + 0.4; labels = torch.randint(0, 50, (512,), generator=g).cuda() % 3
    losses = []
    for _ in range(30):
        opt.zero_grad()
        img, seg = m.render_rays(o, d, 33)
        loss = _head_loss(img, seg, target, labels)
        loss.backward(); opt.step(); losses.append(loss.item())
    print("HEAD training loss", losses[0], "->", losses[-1])
    assert np.isfinite(losses).all() and losses[-1] < 0.7 * losses[0]


def test_head_backward_abi_edges():
    """Status codes of the backward entry points: zero rays are no-ops, bad strides / missing forward outputs /
    short workspaces are refused."""
    from nerf_b200._lib import lib
    L = lib()
    x = torch.zeros(64 * 64, device="cuda")
    assert L.nerf_head_composite_bwd(None, 54, 50, None, None, 0.0, 0, 15, None, None, None, None, 64, None) == 0
    assert L.nerf_head_composite_bwd(x.data_ptr(), 54, 50, x.data_ptr(), None, 0.0, 4, 15, None, None, x.data_ptr(), x.data_ptr(), 64, None) == -1
    assert L.nerf_head_composite_bwd(x.data_ptr(), 54, 50, x.data_ptr(), None, 0.0, 4, 15, None, None, None, x.data_ptr(), 50, None) == -1
    assert L.nerf_head_composite_bwd(x.data_ptr(), 54, 50, x.data_ptr(), None, 0.0, 1, 4000, None, None, None, x.data_ptr(), 64, None) == -2
    assert L.nerf_head_mlp_train_workspace_bytes(0) == 0 and L.nerf_head_mlp_train_workspace_bytes(64) == (7 * 64 * 256 + 10 * 64) * 4
    import ctypes
    pp = (ctypes.c_void_p * 22)(*[x.data_ptr()] * 22)
    assert L.nerf_head_mlp_bwd(pp, x.data_ptr(), 0, 96, 54, x.data_ptr(), 64, pp, None, 0, None) == 0
    assert L.nerf_head_mlp_bwd(pp, x.data_ptr(), 8, 96, 54, x.data_ptr(), 56, pp, x.data_ptr(), 1 << 20, None) == -1
    assert L.nerf_head_mlp_bwd(pp, x.data_ptr(), 8, 96, 54, x.data_ptr(), 64, pp, x.data_ptr(), 16, None) == -3
    assert L.nerf_head_mlp_fwd_train(pp, x.data_ptr(), 8, 96, 54, x.data_ptr(), x.data_ptr(), 16, None) == -3


def test_head_rng_follows_torch_like_the_reference(head_model):
    """nerf/model.py:432,652: every jittered call draws a fresh torch.rand (so two calls differ and torch.manual_seed
    reproduces them), torch.randn is consumed even with density_noise_std == 0, `u` is ignored unless randomly_sample,
    and a jittered render_image does not depend on chunking or on shard boundaries."""
    g = torch.Generator().manual_seed(41)
    o = (torch.randn(50, 3, generator=g) * 2).cuda(); d = torch.randn(50, 3, generator=g).cuda()
    with torch.no_grad():
        torch.manual_seed(7)
        a, _ = head_model.render_rays(o, d, 32, randomly_sample=True)
        b, _ = head_model.render_rays(o, d, 32, randomly_sample=True)
        torch.manual_seed(7)
        a2, _ = head_model.render_rays(o, d, 32, randomly_sample=True)
        assert not torch.equal(a, b) and torch.equal(a, a2)
        # RNG consumption of one deterministic call = rand is NOT drawn, randn IS (reference :652-654)
        torch.manual_seed(9); head_model.render_rays(o, d, 32); after = torch.rand(1, device="cuda")
        torch.manual_seed(9); torch.randn(50, 31, device="cuda"); want = torch.rand(1, device="cuda")
        assert torch.equal(after, want)
        u = torch.rand(50, 32, device="cuda")
        t0 = head_model.sample_along_rays(o, d, 32, randomly_sample=False)
        assert torch.equal(head_model.sample_along_rays(o, d, 32, randomly_sample=False, u=u), t0)
        assert not torch.equal(head_model.sample_along_rays(o, d, 32, randomly_sample=True, u=u), t0)
        cam_o = torch.tensor([[0.0, 0.0, 4.0]], device="cuda"); cam_r = torch.eye(3, device="cuda")[None]
        kw = dict(randomly_sample=True, seed=123)
        full, _ = head_model.render_image(cam_o, cam_r, 20, 30, 40.0, 24, **kw)
        head_model.max_chunk_rays, keep = 100, head_model.max_chunk_rays
        try:
            small, _ = head_model.render_image(cam_o, cam_r, 20, 30, 40.0, 24, max_chunk_size=64, **kw)
            part, _ = head_model.render_image(cam_o, cam_r, 20, 30, 40.0, 24, ray_begin=150, ray_end=420, **kw)
        finally:
            head_model.max_chunk_rays = keep
        assert torch.equal(small, full) and torch.equal(part, full.reshape(-1, 3)[150:420])
