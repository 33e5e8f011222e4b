"""fp32-grade tensor-core forward (`precision="bf16x3"`, csrc/fused_tc_x3.cu): every GEMM operand split into two bf16
parts, three tcgen05 products per term, fp32 accumulation.  The bar is the north star's FP32-mode bar -- composited
pixels within 1e-3 max-abs of the oracle -- on the tensor cores (the reference's arithmetic is fp32 addmm,
/root/reference/nerf/model.py:525-542)."""
import numpy as np
import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu


def _cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.fixture(scope="module")
def g1_ref():
    return load_golden("g1_ref.npz")


@pytest.fixture(scope="module")
def model_x3(nerf_pth_state_dict):
    from nerf_b200.model import NeRF
    m = NeRF(normalize_position=6.0, precision="bf16x3").cuda()
    m.load_state_dict(nerf_pth_state_dict)
    return m


def test_x3_network_outputs_and_pixels_vs_oracle(model_x3, g1_oracle, g1_ref):
    o, d, u = _cuda(g1_ref["kat_o"]), _cuda(g1_ref["kat_d"]), _cuda(g1_ref["kat_u"])
    with torch.no_grad():
        r = model_x3.render_rays_full(o, d, 2.0, 6.0, 64, u=u)
        want = g1_oracle.render_rays_full(o.cpu(), d.cpu(), 2.0, 6.0, 64, u=u.cpu())
    assert np.array_equal(r["t"].cpu().numpy(), g1_ref["kat_t"])                        # fused sampler: bit-exact bins
    es = (r["density"].cpu() - want["density"]).abs() / (1.0 + want["density"].abs())
    ec = (r["color"].cpu() - want["color"]).abs() / (1.0 + want["color"].abs())
    ep = (r["pixels"].cpu() - want["pixels"]).abs()
    print(f"bf16x3 vs oracle: density rel err max {es.max().item():.2e}, colour {ec.max().item():.2e}, pixels max {ep.max().item():.2e}")
    assert es.max().item() <= 2e-3 and ec.max().item() <= 2e-3
    assert ep.max().item() <= 1e-3                                                     # the fp32-mode bar of the north star


def test_x3_render_image_d1_fp32_bar(model_x3, g1_ref):
    """D1 (100 x 100 x 64, the notebook's held-out pose): pixels <= 1e-3 max-abs vs the oracle image; the camera path,
    the explicit-ray path and a ray shard agree bit for bit."""
    from oracle import g1
    from nerf_b200 import ops
    pose = g1.pose_spherical(*g1.FIXTURE_POSE).cuda()
    co, cr = pose[None, :3, 3].contiguous(), pose[None, :3, :3].contiguous()
    with torch.no_grad():
        img = model_x3.render_image(co, cr, 100, 100, g1.TINY_NERF_FOCAL, 2.0, 6.0, 64)
        part = model_x3.render_image(co, cr, 100, 100, g1.TINY_NERF_FOCAL, 2.0, 6.0, 64, ray_begin=3001, ray_end=7777)
        o, d = ops.generate_rays(co, cr, 100, 100, g1.TINY_NERF_FOCAL)
        explicit = model_x3.render_rays_full(o, d, 2.0, 6.0, 64)["pixels"]
    err = np.abs(img[0].cpu().numpy() - g1_ref["d1_fixture_image"])
    print(f"bf16x3 D1 image vs oracle: max|diff| {err.max():.2e}, mean {err.mean():.2e}")
    assert err.max() <= 1e-3
    assert torch.equal(part, img.reshape(-1, 3)[3001:7777]) and torch.equal(explicit, img.reshape(-1, 3))
    assert ops.COUNTERS["g1_render_fused_x3"] >= 3


@pytest.mark.parametrize("b,s", [(1, 32), (3, 40), (7, 64), (5, 192), (300, 64), (129, 33)])
def test_x3_ragged_shapes_vs_fp32(model_x3, nerf_pth_state_dict, b, s):
    """rays straddling tiles, a partial last tile, an odd number of tiles (the CTA pair's second tile is empty)"""
    from nerf_b200.model import NeRF
    m32 = NeRF(normalize_position=6.0, precision="fp32").cuda(); m32.load_state_dict(nerf_pth_state_dict)
    gen = torch.Generator().manual_seed(b * 1000 + s)
    o = torch.zeros(b, 3); o[:, 2] = 4.0
    d = torch.randn(b, 3, generator=gen) * 0.15; d[:, 2] = -1.0
    with torch.no_grad():
        a = model_x3.render_rays_full(o.cuda(), d.cuda(), 2.0, 6.0, s)
        w = m32.render_rays_full(o.cuda(), d.cuda(), 2.0, 6.0, s)
    assert torch.equal(a["t"], w["t"]) and torch.isfinite(a["pixels"]).all()
    err = (a["pixels"] - w["pixels"]).abs().max().item()
    assert err <= 1e-3, err


def test_x3_hierarchical_and_throughput(model_x3, nerf_pth_state_dict, g1_ref):
    """the hierarchical pass on the x3 forward against the oracle's pixels, and the point of the mode: at least 5x the
    fp32 FFMA path's render throughput at fp32-grade accuracy"""
    from oracle import g1
    from nerf_b200.model import NeRF
    with torch.no_grad():
        r = model_x3.render_rays_hierarchical(_cuda(g1_ref["kat_o"]), _cuda(g1_ref["kat_d"]), 2.0, 6.0, 64, 128)
    err = np.abs(r["pixels"].cpu().numpy() - g1_ref["kat_hier_pixels"]).max()
    print(f"bf16x3 hierarchical pixels vs oracle: max|diff| {err:.2e}")
    assert err <= 2e-3
    m32 = NeRF(normalize_position=6.0, precision="fp32").cuda(); m32.load_state_dict(nerf_pth_state_dict)
    pose = g1.pose_spherical(30.0, -30.0, 4.0311).cuda()
    co, cr = pose[None, :3, 3].contiguous(), pose[None, :3, :3].contiguous()

    def rays_per_s(m, hw, n=2):
        with torch.no_grad():
            m.render_image(co, cr, hw, hw, g1.TINY_NERF_FOCAL * hw / 100, 2.0, 6.0, 64)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                m.render_image(co, cr, hw, hw, g1.TINY_NERF_FOCAL * hw / 100, 2.0, 6.0, 64)
            e1.record(); torch.cuda.synchronize()
        return n * hw * hw / (e0.elapsed_time(e1) * 1e-3)
    x3, f32 = rays_per_s(model_x3, 800), rays_per_s(m32, 200)
    print(f"render throughput: bf16x3 (tcgen05, fp32-grade) {x3 / 1e6:.2f} M rays/s = {x3 * 64 * 1261568 / 1e12:.0f} TFLOP/s; "
          f"fp32 FFMA {f32 / 1e6:.2f} M rays/s = {f32 * 64 * 1261568 / 1e12:.0f} TFLOP/s; ratio {x3 / f32:.1f}x")
    assert x3 >= 5.0 * f32
