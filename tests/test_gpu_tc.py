"""tcgen05 (bf16) path against the oracle: per-unit activations, sigma/rgb, pixels and the
north-star bar "PSNR within 0.05 dB in bf16 mode"."""
import numpy as np
import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu


def _cuda(a):
    return torch.from_numpy(np.asarray(a)).cuda()


@pytest.fixture(scope="module")
def model_bf16(nerf_pth_state_dict):
    from nerf_b200.model import NeRF
    m = NeRF(normalize_position=6.0, precision="bf16").cuda()
    m.load_state_dict(nerf_pth_state_dict, strict=True)
    return m


@pytest.fixture(scope="module")
def packed(model_bf16):
    from nerf_b200 import ops
    return ops.g1_pack_weights(model_bf16.flat_parameters())


def _oracle_acts(g1_oracle, o, d, t):
    with torch.no_grad():
        _, ex, ed = g1_oracle.encode(o, d, t)
        density, color, acts = g1_oracle.mlp(ex, ed, return_activations=True)
    return density, color, acts


@pytest.mark.parametrize("layer", [0, 1, 3, 4, 7, 8, 9])
def test_tc_unit_activations(g1_oracle, g1_ref, packed, layer):
    """Normalised output of one unit (before bf16 rounding) with the unit's affine applied on
    the host, against the oracle's activation.  Errors come from bf16 operands only."""
    from nerf_b200 import ops
    o, d, t = _cuda(g1_ref["kat_o"]), _cuda(g1_ref["kat_d"]), _cuda(g1_ref["kat_t"])
    _, _, dbg = ops.g1_mlp_tc_debug(packed, o, d, t, 6.0, dbg_layer=layer)
    _, _, acts = _oracle_acts(g1_oracle, o.cpu(), d.cpu(), t.cpu())
    blocks = [g1_oracle.block_0, g1_oracle.block_0, g1_oracle.block_0, g1_oracle.block_0,
              g1_oracle.block_1, g1_oracle.block_1, g1_oracle.block_1, g1_oracle.block_1,
              g1_oracle.block_2, g1_oracle.block_2]
    ln = blocks[layer][3 * (layer % 4 if layer < 8 else layer - 8) + 2]
    got = dbg.cpu() * ln.weight.detach() + ln.bias.detach()
    want = acts[layer].reshape(-1, 256)
    err = (got - want).abs()
    rel = (got - want).norm() / want.norm()
    print(f"unit {layer}: max abs {err.max().item():.4f}, rel-L2 {rel.item():.5f}")
    assert rel.item() < 0.03 * (1 + layer // 4), (layer, rel.item())


def test_tc_sigma_rgb_and_pixels(g1_oracle, g1_ref, packed, model_bf16):
    from nerf_b200 import ops
    o, d, t = _cuda(g1_ref["kat_o"]), _cuda(g1_ref["kat_d"]), _cuda(g1_ref["kat_t"])
    sigma, rgb, _ = ops.g1_mlp_tc_debug(packed, o, d, t, 6.0)
    wsig, wrgb, _ = _oracle_acts(g1_oracle, o.cpu(), d.cpu(), t.cpu())
    es = (sigma.cpu() - wsig[..., 0]).abs(); ec = (rgb.cpu() - wrgb).abs()
    print("sigma max/mean err", es.max().item(), es.mean().item(), "rgb", ec.max().item(), ec.mean().item())
    assert es.mean().item() < 0.15 and ec.mean().item() < 0.05
    with torch.no_grad():
        r = model_bf16.render_rays_full(o, d, 2.0, 6.0, 64, u=_cuda(g1_ref["kat_u"]))
        want = g1_oracle.render_rays_full(o.cpu(), d.cpu(), 2.0, 6.0, 64, u=torch.from_numpy(g1_ref["kat_u"]))
    assert np.array_equal(r["t"].cpu().numpy(), g1_ref["kat_t"])            # fused sampler: bit-exact bins
    perr = (r["pixels"].cpu() - want["pixels"]).abs()
    print("pixel max/mean err (bf16)", perr.max().item(), perr.mean().item())
    assert perr.max().item() < 0.05 and perr.mean().item() < 5e-3


def test_tc_render_image_psnr(model_bf16, g1_ref):
    """D1 in bf16 mode: PSNR against the notebook's ground-truth panel within 0.05 dB of the
    fp32 oracle's, and the camera path == explicit-ray path == any ray shard, bit for bit."""
    from oracle import g1
    from nerf_b200 import ops
    pose = g1.pose_spherical(*g1.FIXTURE_POSE).cuda()
    co, cr = pose[None, :3, 3].contiguous(), pose[None, :3, :3].contiguous()
    with torch.no_grad():
        img = model_bf16.render_image(co, cr, 100, 100, g1.TINY_NERF_FOCAL, 2.0, 6.0, 64)
        part = model_bf16.render_image(co, cr, 100, 100, g1.TINY_NERF_FOCAL, 2.0, 6.0, 64, ray_begin=3001, ray_end=7777)
        o, d = ops.generate_rays(co, cr, 100, 100, g1.TINY_NERF_FOCAL)
        explicit = model_bf16.render_rays_full(o, d, 2.0, 6.0, 64)["pixels"]
    got = img[0].cpu().numpy()
    panel = load_golden("notebook_panel.npz")
    psnr = lambda a, b: -10 * np.log10(((a - b) ** 2).mean())
    p_bf16, p_oracle = psnr(got, panel["truth"]), psnr(g1_ref["d1_fixture_image"], panel["truth"])
    p_vs_oracle = psnr(got, g1_ref["d1_fixture_image"])
    print(f"PSNR vs notebook truth: bf16 {p_bf16:.3f} dB, oracle {p_oracle:.3f} dB; bf16 vs oracle image {p_vs_oracle:.2f} dB")
    assert abs(p_bf16 - p_oracle) <= 0.05
    assert psnr(got, panel["render"]) >= 32.0
    assert torch.equal(part, img.reshape(-1, 3)[3001:7777])
    assert torch.equal(explicit, img.reshape(-1, 3))


@pytest.mark.parametrize("b,s", [(1, 32), (3, 40), (7, 64), (5, 192), (300, 64)])
def test_tc_ragged_shapes_vs_fp32(model_bf16, nerf_pth_state_dict, b, s):
    """Tiles that straddle rays / partial last tile: bf16 fused path vs our fp32 kernels."""
    from nerf_b200.model import NeRF
    m32 = NeRF(normalize_position=6.0).cuda(); m32.load_state_dict(nerf_pth_state_dict)
    gen = torch.Generator().manual_seed(b * 1000 + s)
    o = torch.zeros(b, 3); o[:, 2] = 4.0
    d = torch.randn(b, 3, generator=gen) * 0.15; d[:, 2] = -1.0
    with torch.no_grad():
        a = model_bf16.render_rays_full(o.cuda(), d.cuda(), 2.0, 6.0, s)
        w = m32.render_rays_full(o.cuda(), d.cuda(), 2.0, 6.0, s)
    assert torch.equal(a["t"], w["t"])
    err = (a["pixels"] - w["pixels"]).abs()
    assert torch.isfinite(a["pixels"]).all()
    assert err.max().item() < 0.06, err.max().item()


def test_tc_hierarchical(model_bf16, g1_ref):
    with torch.no_grad():
        r = model_bf16.render_rays_hierarchical(_cuda(g1_ref["kat_o"]), _cuda(g1_ref["kat_d"]), 2.0, 6.0, 64, 128)
    err = np.abs(r["pixels"].cpu().numpy() - g1_ref["kat_hier_pixels"])
    print("hierarchical bf16 pixel err max/mean", err.max(), err.mean())
    assert err.max() < 0.06 and err.mean() < 6e-3
