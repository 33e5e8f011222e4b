"""Regenerates tests/golden/head_grad_ref.npz: loss and parameter gradients of the UNMODIFIED reference
nerf/model.py (HEAD generation, imported from /root/reference) through torch autograd.  Build container
only (the GPU box has no /root/reference).  Usage:  python tests/golden/make_head_grad.py

Two cases on the model and rays of head_ref.npz (torch.manual_seed(3) default init):
  det    render_rays(ro, rd, 16, randomly_sample=False, density_noise_std=0.0)
  noisy  torch.manual_seed(9); render_rays(ro, rd, 16, randomly_sample=True, density_noise_std=1.0)
         (u = the reference's own torch.rand draw, noise = its torch.randn draw; both stored)
loss = mean((image - target)^2) + nll_loss(segmentation, labels): colour and segmentation branches both live.
"""
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def main():
    ref = np.load(os.path.join(HERE, "head_ref.npz"))
    sys.path.insert(0, REF)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        from nerf.model import NeRF as RefNeRF
        torch.manual_seed(3)
        m = RefNeRF()
        for k, v in m.state_dict().items():
            assert np.array_equal(v.numpy(), ref["head_sd/" + k]), k          # same model as head_ref.npz
        ro, rd = torch.from_numpy(ref["head_ro"]), torch.from_numpy(ref["head_rd"])
        g = torch.Generator().manual_seed(31)
        target = torch.rand(6, 3, generator=g)
        labels = torch.randint(0, 50, (6,), generator=g)
        out = dict(target=target.numpy(), labels=labels.numpy())

        def run(tag, **kw):
            m.zero_grad()
            img, seg = m.render_rays(ro, rd, 16, **kw)
            loss = ((img[:, 0] - target) ** 2).mean() + torch.nn.functional.nll_loss(seg[:, 0], labels)
            loss.backward()
            out[tag + "_loss"] = loss.detach().numpy()
            out[tag + "_img"] = img.detach().numpy(); out[tag + "_seg"] = seg.detach().numpy()
            for k, p in m.named_parameters():
                out[f"{tag}_grad/{k}"] = p.grad.numpy().copy()

        run("det", randomly_sample=False, density_noise_std=0.0)
        torch.manual_seed(9)
        out["noisy_u"] = torch.rand(6, 16).numpy()
        out["noisy_noise"] = torch.randn(6, 15, 1).numpy()
        torch.manual_seed(9)
        run("noisy", randomly_sample=True, density_noise_std=1.0)
    path = os.path.join(HERE, "head_grad_ref.npz")
    np.savez_compressed(path, **out)
    print(f"head_grad_ref.npz: {os.path.getsize(path) / 1024:.1f} KiB")


if __name__ == "__main__":
    main()
