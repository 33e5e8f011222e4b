"""Regenerates tests/golden/*.npz.  Runs ONLY in the build container (it reads
/root/reference, which does not exist on the GPU box).  Usage:

    python tests/golden/make_golden.py

Outputs
-------
nerf_pth.npz        the 44 fp32 tensors of /root/reference/examples/nerf.pth, key
                    order preserved (``__keys__``), so the GPU box can rebuild the
                    checkpoint byte-for-byte in meaning.
notebook_panel.npz  the "Iteration: 40000" render and "Ground Truth" panels
                    cropped from example.ipynb cell 8 output #40,
                    area-downsampled to 100x100 float32.
head_ref.npz        outputs of the UNMODIFIED reference nerf/model.py (HEAD),
                    imported from /root/reference, on small seeded inputs:
                    generate_rays, rays_to_world_coordinates,
                    alpha_compositing_coefficients, sample_along_rays,
                    integrated_pe, forward, render_rays, render_image.
g1_ref.npz          outputs of oracle.g1 with nerf.pth on the D1/D2 inputs
                    (regression pins for the GPU box; oracle itself is pinned by
                    the notebook panel + head_ref shared functions).
"""
import base64
import io
import json
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)


def save(name, **arrays):
    path = os.path.join(HERE, name)
    np.savez_compressed(path, **arrays)
    print(f"{name}: {os.path.getsize(path) / 1024:.1f} KiB")


def make_nerf_pth():
    sd = torch.load(os.path.join(REF, "examples/nerf.pth"), map_location="cpu")
    arrays = {k: v.numpy() for k, v in sd.items()}
    save("nerf_pth.npz", __keys__=np.array(list(sd.keys())), **arrays)
    return sd


def make_panel():
    import cv2
    nb = json.load(open(os.path.join(REF, "examples/example.ipynb")))
    png = base64.b64decode(nb["cells"][8]["outputs"][40]["data"]["image/png"])
    img = cv2.imdecode(np.frombuffer(png, np.uint8), cv2.IMREAD_UNCHANGED)
    img = cv2.cvtColor(img, cv2.COLOR_BGRA2RGB).astype(np.float32) / 255.0
    render = img[23:240, 28:244]
    truth = img[23:240, 323:540]
    small = lambda a: cv2.resize(a, (100, 100), interpolation=cv2.INTER_AREA)
    save("notebook_panel.npz", render=small(render), truth=small(truth))


def make_head_ref():
    sys.path.insert(0, REF)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        from nerf.model import NeRF as RefNeRF
        import nerf.model as refmod
        out = {}
        for (h, w, f) in [(5, 7, 3.3), (100, 100, 138.88887889922103), (3, 2, 50.0)]:
            out[f"rays_{h}_{w}"] = RefNeRF.generate_rays(h, w, f).numpy()
        g = torch.Generator().manual_seed(11)
        rays = torch.randn(4, 6, 3, generator=g)
        cam_o = torch.randn(4, 6, 3, generator=g)
        cam_r = torch.randn(4, 6, 3, 3, generator=g)
        o, d = RefNeRF.rays_to_world_coordinates(rays, cam_o, cam_r)
        out.update(w2c_rays=rays.numpy(), w2c_cam_o=cam_o.numpy(), w2c_cam_r=cam_r.numpy(),
                   w2c_o=o.numpy(), w2c_d=d.numpy())
        # compositing weights on random + adversarial inputs
        pts = torch.cumsum(torch.rand(9, 17, 3, generator=g) * 0.1, dim=1)
        sig = torch.randn(9, 17, 1, generator=g) * 20.0
        sig[0] = -1.0          # empty ray
        sig[1] = 1e4           # opaque at first sample
        out.update(comp_points=pts.numpy(), comp_density=sig.numpy(),
                   comp_weights=RefNeRF.alpha_compositing_coefficients(pts, sig).numpy())
        # HEAD model pieces with seeded default init
        torch.manual_seed(3)
        m = RefNeRF()
        sd = m.state_dict()
        out["head_keys"] = np.array(list(sd.keys()))
        for k, v in sd.items():
            out["head_sd/" + k] = v.numpy()
        ro = torch.randn(6, 3, generator=g) * 2
        rd = torch.randn(6, 3, generator=g)
        out.update(head_ro=ro.numpy(), head_rd=rd.numpy())
        out["head_t_det"] = m.sample_along_rays(ro, rd, 16, randomly_sample=False).numpy()
        torch.manual_seed(5)
        out["head_t_rand"] = m.sample_along_rays(ro, rd, 16, randomly_sample=True).numpy()
        torch.manual_seed(5)
        out["head_u"] = torch.rand(6, 16).numpy()
        t = torch.from_numpy(out["head_t_rand"])
        mean, cov, feat = m.integrated_pe(ro, rd, t)
        out.update(head_ipe_mean=mean.numpy(), head_ipe_cov=cov.numpy(), head_ipe_feat=feat.numpy())
        mean, dens, col, seg = m.forward(ro, rd, t)
        out.update(head_fwd_density=dens.detach().numpy(), head_fwd_color=col.detach().numpy(),
                   head_fwd_seg=seg.detach().numpy())
        torch.manual_seed(7)
        img, segm = m.render_rays(ro, rd, 16, randomly_sample=False, density_noise_std=0.0)
        out.update(head_rr_img=img.detach().numpy(), head_rr_seg=segm.detach().numpy())
        co = torch.tensor([[0.0, 0.0, 4.0], [1.0, -2.0, 3.0]])
        cr = torch.linalg.qr(torch.randn(2, 3, 3, generator=g))[0]
        img, segm = m.render_image(co, cr, 6, 5, 7.5, 12, max_chunk_size=7)
        out.update(head_ri_co=co.numpy(), head_ri_cr=cr.numpy(),
                   head_ri_img=img.detach().numpy(), head_ri_seg=segm.detach().numpy())
    sys.path.remove(REF)
    for k in [k for k in sys.modules if k == "nerf" or k.startswith("nerf.")]:
        del sys.modules[k]
    save("head_ref.npz", **out)


def make_g1_ref(sd):
    from oracle import g1
    m = g1.G1NeRF(normalize_position=6.0)
    m.load_state_dict(sd, strict=True)
    m.eval()
    out = {}
    with torch.no_grad():
        pose = g1.pose_spherical(*g1.FIXTURE_POSE)
        img = m.render_image(pose[None, :3, 3], pose[None, :3, :3], 100, 100,
                             g1.TINY_NERF_FOCAL, 2.0, 6.0, 64)
        out["d1_fixture_image"] = img[0].numpy()
        # a 32x32 crop of a second pose, full intermediates for kernel-level checks
        pose2 = g1.pose_spherical(45.0, -30.0, 4.0311)
        rays = g1.generate_rays(100, 100, g1.TINY_NERF_FOCAL)[34:66:4, 34:66:4].reshape(-1, 3)
        o, d = g1.rays_to_world_coordinates(rays, pose2[:3, 3].expand_as(rays),
                                            pose2[:3, :3].expand(rays.shape[0], 3, 3))
        gen = torch.Generator().manual_seed(1)
        u = torch.rand(rays.shape[0], 64, generator=gen)
        noise = torch.randn(rays.shape[0], 64, 1, generator=torch.Generator().manual_seed(2))
        r = m.render_rays_full(o, d, 2.0, 6.0, 64, u=u, noise=noise)
        out.update(kat_o=o.numpy(), kat_d=d.numpy(), kat_u=u.numpy(), kat_noise=noise.numpy(),
                   kat_t=r["t"].numpy(), kat_density=r["density"].numpy(),
                   kat_color=r["color"].numpy(), kat_weights=r["weights"].numpy(),
                   kat_pixels=r["pixels"].numpy(), kat_depth=r["depth"].numpy(),
                   kat_acc=r["acc"].numpy())
        pts, ex, ed = m.encode(o, d, r["t"])
        out.update(kat_ex=ex[:8].numpy(), kat_ed=ed.numpy())
        h = m.render_rays_hierarchical(o, d, 2.0, 6.0, 64, 128)
        out.update(kat_hier_t=h["t"].numpy(), kat_hier_pixels=h["pixels"].numpy())
    # gradients of the D2-style loss on this small batch (all 44 tensors)
    m.train()
    r = m.render_rays_full(o, d, 2.0, 6.0, 64, u=u, noise=noise)
    target = torch.from_numpy(out["kat_pixels"]) * 0.5 + 0.1
    loss = ((r["pixels"] - target) ** 2).mean()
    loss.backward()
    out["kat_target"] = target.numpy()
    out["kat_loss"] = loss.detach().numpy()
    for k, p in m.named_parameters():
        out["kat_grad/" + k] = p.grad.numpy()
    save("g1_ref.npz", **out)


if __name__ == "__main__":
    sd = make_nerf_pth()
    make_panel()
    make_head_ref()
    make_g1_ref(sd)
