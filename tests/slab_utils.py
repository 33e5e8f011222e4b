"""Host-side encode/decode of the slab image (include/nerf_b200.h, "Slab image") — test tooling
only: the product path never converts, its kernels read and write slabs directly."""
import torch


def _perm(device):
    r = torch.arange(128, device=device)[:, None]
    c = torch.arange(8, device=device)[None, :]
    return (c ^ (r & 7))                      # physical chunk of logical chunk c in row r


def to_slabs(x):
    """x [P, 64*ns] (any float dtype) -> uint8 image [n_tiles, ns, 128, 128] of bf16 (P padded with zeros)."""
    p, k = x.shape
    ns = k // 64
    nt = (p + 127) // 128
    xb = torch.zeros(nt * 128, k, dtype=torch.bfloat16, device=x.device)
    xb[:p] = x.to(torch.bfloat16)
    v = xb.view(nt, 128, ns, 8, 8).permute(0, 2, 1, 3, 4).contiguous()      # [nt, ns, row, chunk, 8]
    out = torch.empty_like(v)
    idx = _perm(x.device)[None, None, :, :, None].expand_as(v)
    out.scatter_(3, idx, v)
    return out.view(torch.uint8).view(nt, ns, 128, 128)


def from_slabs(img, p=None):
    """inverse of to_slabs: uint8 [n_tiles, ns, 128, 128] -> float32 [P, 64*ns]"""
    nt, ns = img.shape[0], img.shape[1]
    v = img.contiguous().view(torch.bfloat16).view(nt, ns, 128, 8, 8)
    idx = _perm(img.device)[None, None, :, :, None].expand_as(v)
    lin = torch.gather(v, 3, idx)
    x = lin.permute(0, 2, 1, 3, 4).reshape(nt * 128, ns * 64).float()
    return x if p is None else x[:p]
