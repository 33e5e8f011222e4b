"""Pins the oracle: HEAD restatement vs the unmodified reference's
stored outputs; G1 restatement vs nerf.pth structure + the notebook panel."""
import os
from collections import OrderedDict

import numpy as np
import torch

from oracle import g1, head
from conftest import load_golden


def _t(a):
    return torch.from_numpy(np.asarray(a))


def test_g1_checkpoint_keys_and_shapes(nerf_pth_state_dict):
    m = g1.G1NeRF(normalize_position=6.0)
    assert list(m.state_dict().keys()) == list(nerf_pth_state_dict.keys())
    assert sum(v.numel() for v in nerf_pth_state_dict.values()) == 638468
    m.load_state_dict(nerf_pth_state_dict, strict=True)


def test_g1_fixture_matches_reference_file_if_present(nerf_pth_state_dict):
    path = "/root/reference/examples/nerf.pth"
    if not os.path.exists(path):
        import pytest
        pytest.skip("reference tree not on this machine")
    sd = torch.load(path, map_location="cpu")
    assert list(sd.keys()) == list(nerf_pth_state_dict.keys())
    for k in sd:
        assert torch.equal(sd[k], nerf_pth_state_dict[k])


def test_g1_reproduces_notebook_render(g1_oracle):
    """>= 32 dB against the embedded 'Iteration: 40000' panel."""
    panel = load_golden("notebook_panel.npz")
    pose = g1.pose_spherical(*g1.FIXTURE_POSE)
    with torch.no_grad():
        img = g1_oracle.render_image(pose[None, :3, 3], pose[None, :3, :3], 100, 100,
                                     g1.TINY_NERF_FOCAL, 2.0, 6.0, 64)[0].numpy()
    psnr = lambda a, b: -10 * np.log10(((a - b) ** 2).mean())
    assert psnr(img, panel["render"]) >= 32.0
    assert psnr(img, panel["truth"]) >= 30.5


def test_g1_regression_vectors(g1_oracle, g1_ref):
    with torch.no_grad():
        r = g1_oracle.render_rays_full(_t(g1_ref["kat_o"]), _t(g1_ref["kat_d"]), 2.0, 6.0, 64,
                                       u=_t(g1_ref["kat_u"]), noise=_t(g1_ref["kat_noise"]))
    np.testing.assert_allclose(r["t"].numpy(), g1_ref["kat_t"], rtol=0, atol=1e-6)
    np.testing.assert_allclose(r["pixels"].numpy(), g1_ref["kat_pixels"], rtol=0, atol=2e-5)


def test_g1_chunk_invariance(g1_oracle):
    pose = g1.pose_spherical(30.0, -30.0, 4.0311)
    with torch.no_grad():
        a = g1_oracle.render_image(pose[None, :3, 3], pose[None, :3, :3], 12, 10, 15.0, 2.0, 6.0, 16, max_chunk_size=7)
        b = g1_oracle.render_image(pose[None, :3, 3], pose[None, :3, :3], 12, 10, 15.0, 2.0, 6.0, 16, max_chunk_size=1024)
    np.testing.assert_allclose(a.numpy(), b.numpy(), atol=2e-6)


# ---- HEAD restatement vs the real reference's stored outputs -----------------
def test_generate_rays_vs_reference(head_ref):
    for (h, w, f) in [(5, 7, 3.3), (100, 100, 138.88887889922103), (3, 2, 50.0)]:
        got = g1.generate_rays(h, w, f).numpy()
        assert np.array_equal(got, head_ref[f"rays_{h}_{w}"])


def test_world_transform_vs_reference(head_ref):
    o, d = g1.rays_to_world_coordinates(_t(head_ref["w2c_rays"]), _t(head_ref["w2c_cam_o"]),
                                        _t(head_ref["w2c_cam_r"]))
    assert np.array_equal(d.numpy(), head_ref["w2c_d"])
    assert np.array_equal(o.numpy(), head_ref["w2c_o"])


def test_compositing_vs_reference(head_ref):
    w = g1.alpha_compositing_coefficients(_t(head_ref["comp_points"]), _t(head_ref["comp_density"]))
    assert np.array_equal(w.numpy(), head_ref["comp_weights"])


def _head_model(head_ref):
    m = head.HeadNeRF()
    sd = OrderedDict((str(k), _t(head_ref["head_sd/" + str(k)])) for k in head_ref["head_keys"])
    assert list(m.state_dict().keys()) == list(sd.keys())
    m.load_state_dict(sd, strict=True)
    return m.eval()


def test_head_sampler_vs_reference(head_ref):
    m = _head_model(head_ref)
    t = head.sample_along_rays([6], 16, m.box_diag())
    assert np.array_equal(t.numpy(), head_ref["head_t_det"])
    t = head.sample_along_rays([6], 16, m.box_diag(), u=_t(head_ref["head_u"]))
    assert np.array_equal(t.numpy(), head_ref["head_t_rand"])


def test_head_encoder_and_trunk_vs_reference(head_ref):
    m = _head_model(head_ref)
    ro, rd, t = _t(head_ref["head_ro"]), _t(head_ref["head_rd"]), _t(head_ref["head_t_rand"])
    mean, cov, feat = m.encode(ro, rd, t)
    np.testing.assert_allclose(mean.numpy(), head_ref["head_ipe_mean"], rtol=0, atol=0)
    np.testing.assert_allclose(cov.numpy(), head_ref["head_ipe_cov"], rtol=1e-6, atol=0)
    np.testing.assert_allclose(feat.numpy(), head_ref["head_ipe_feat"], rtol=0, atol=1e-6)
    with torch.no_grad():
        _, dens, col, seg = m.forward(ro, rd, t)
    np.testing.assert_allclose(dens.numpy(), head_ref["head_fwd_density"], atol=1e-5)
    np.testing.assert_allclose(col.numpy(), head_ref["head_fwd_color"], atol=1e-5)
    np.testing.assert_allclose(seg.numpy(), head_ref["head_fwd_seg"], atol=1e-5)


def test_head_render_vs_reference(head_ref):
    m = _head_model(head_ref)
    with torch.no_grad():
        img, seg = m.render_rays(_t(head_ref["head_ro"]), _t(head_ref["head_rd"]), 16)
        np.testing.assert_allclose(img.numpy(), head_ref["head_rr_img"], atol=1e-6)
        np.testing.assert_allclose(seg.numpy(), head_ref["head_rr_seg"], atol=1e-4)
        img, seg = m.render_image(_t(head_ref["head_ri_co"]), _t(head_ref["head_ri_cr"]), 6, 5, 7.5, 12,
                                  max_chunk_size=7)
        np.testing.assert_allclose(img.numpy(), head_ref["head_ri_img"], atol=1e-6)
        np.testing.assert_allclose(seg.numpy(), head_ref["head_ri_seg"], atol=1e-4)


def test_head_gradients_vs_reference(head_ref):
    """Autograd through the HEAD oracle against the UNMODIFIED reference's parameter gradients
    (tests/golden/head_grad_ref.npz, tests/golden/make_head_grad.py): pins the checker the CUDA backward is
    compared with at sizes the fixture does not hold."""
    from oracle import head as oh
    gr = load_golden("head_grad_ref.npz")
    m = oh.HeadNeRF()
    m.load_state_dict({str(k): torch.from_numpy(head_ref["head_sd/" + str(k)]) for k in head_ref["head_keys"]})
    ro, rd = torch.from_numpy(head_ref["head_ro"]), torch.from_numpy(head_ref["head_rd"])
    target, labels = torch.from_numpy(gr["target"]), torch.from_numpy(gr["labels"])
    for tag, kw in (("det", {}), ("noisy", dict(u=torch.from_numpy(gr["noisy_u"]), noise=torch.from_numpy(gr["noisy_noise"])))):
        m.zero_grad()
        img, seg = m.render_rays(ro, rd, 16, **kw)
        loss = ((img[:, 0] - target) ** 2).mean() + torch.nn.functional.nll_loss(seg[:, 0], labels)
        loss.backward()
        np.testing.assert_allclose(loss.item(), gr[tag + "_loss"], rtol=1e-6)
        for k, p in m.named_parameters():
            want = gr[f"{tag}_grad/{k}"]
            np.testing.assert_allclose(p.grad.numpy(), want, rtol=1e-4, atol=1e-6 * max(1.0, np.abs(want).max()), err_msg=f"{tag} {k}")


def test_hierarchical_sampler_known_answers():
    """oracle.g1.sample_pdf against the hand-derived vectors of tests/golden/make_pdf_kat.py (NeRF section 5.2; the
    algorithm the reference promises at nerf/model.py:191-193, 503-505 and never implements)."""
    from conftest import load_golden
    from oracle import g1
    kat = load_golden("pdf_kat.npz")
    for name in kat["names"]:
        t = torch.from_numpy(kat[f"{name}_t"])[None]; w = torch.from_numpy(kat[f"{name}_w"])[None]
        u = kat[f"{name}_u"]
        want = kat[f"{name}_fine"]
        fine = g1.sample_pdf(t, w, len(want), u=None if u.size == 0 else torch.from_numpy(u)[None])[0].numpy()
        np.testing.assert_allclose(fine, want, rtol=0, atol=1e-5 if name == "C" else 5e-6, err_msg=str(name))
        merged = np.sort(np.concatenate([t[0].numpy(), fine]))
        np.testing.assert_allclose(merged, kat[f"{name}_merged"], rtol=0, atol=1e-5)
