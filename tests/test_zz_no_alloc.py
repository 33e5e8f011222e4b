"""`nerf_g1_render_fused_static`: the fused render with its scratch in a file-scope device array of the library, for
callers that cannot allocate (CUDA graph capture, serving loops).  Called straight through the C ABI and compared with
the existing entry point `nerf_g1_render_fused` (caller workspace) on consecutive calls that share the array.

Collected last on purpose; passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(b, seed):
    g = torch.Generator().manual_seed(seed)
    o = torch.zeros(b, 3); o[:, 2] = 4.0
    o[:, :2] = torch.randn(b, 2, generator=g) * 0.2
    d = torch.randn(b, 3, generator=g) * 0.2; d[:, 2] = -1.0
    return o.cuda().contiguous(), d.cuda().contiguous()


def _render_both(L, P, st, packed, ladder, s, rays=None, camera=None, ray_begin=0, b=None, u=None):
    """-> (pixels, depth, acc) of the existing entry point and of the no-allocation one, same arguments"""
    from nerf_b200 import _lib
    if rays is not None:
        o, d = rays; b = o.shape[0]
        cam = (None, None, 0, 0, 0, 1.0)
    else:
        o = d = None
        cam = camera
    outs = []
    for static in (False, True):
        px = torch.full((b, 3), float("nan"), device="cuda"); dp = torch.full((b,), float("nan"), device="cuda")
        ac = torch.full((b,), float("nan"), device="cuda")
        if static:
            rc = L.nerf_g1_render_fused_static(P(packed), P(o), P(d), P(cam[0]), P(cam[1]), cam[2], cam[3], cam[4], cam[5], ray_begin, b, s,
                                               2.0, 6.0, P(ladder), P(u), 0, 0, 0, 0.0, 6.0, 1, P(px), P(dp), P(ac), st())
        else:
            n = L.nerf_g1_render_workspace_bytes(b, s, 3, 0)
            ws = torch.empty(n, dtype=torch.uint8, device="cuda")
            rc = L.nerf_g1_render_fused(P(packed), 3, 3, P(o), P(d), P(cam[0]), P(cam[1]), cam[2], cam[3], cam[4], cam[5], ray_begin, b, s,
                                        2.0, 6.0, P(ladder), None, P(u), 0, 0, 0, None, 0.0, 6.0, 1, None, 0, 1,
                                        P(px), P(dp), P(ac), None, None, None, None, P(ws), n, st())
        _lib.check(rc, "render (static)" if static else "render")
        torch.cuda.synchronize()
        outs.append((px, dp, ac))
    return outs


def test_static_scratch_render_matches_workspace_render(nerf_pth_state_dict):
    from nerf_b200 import _lib, ops
    from nerf_b200.model import NeRF
    from oracle import g1
    L, P, st = _lib.lib(), _lib.ptr, _lib.stream
    m = NeRF(normalize_position=6.0, precision="bf16").cuda(); m.load_state_dict(nerf_pth_state_dict)
    packed = ops.g1_pack_weights(m.flat_parameters(), 3, 3)
    s = 64
    ladder = torch.linspace(2.0, 6.0, s).cuda()
    assert L.nerf_g1_render_static_max_points() == 131072

    def same(a, b, what):
        for x, y, name in zip(a, b, ("pixels", "depth", "acc")):
            assert torch.isfinite(y).all(), (what, name)
            err = (x - y).abs().max().item()
            assert err <= 1e-6, (what, name, err)           # same kernels on the same inputs: equal; 1e-6 = the fp32 pixel tolerance of the suite

    # call 1: the largest size the array takes (2048 rays x 64 = 131072 points), jittered with host uniforms
    o1, d1 = _inputs(2048, 1)
    u1 = torch.rand(2048, s, generator=torch.Generator().manual_seed(2)).cuda()
    ref1, got1 = _render_both(L, P, st, packed, ladder, s, rays=(o1, d1), u=u1)
    same(ref1, got1, "call 1")
    assert got1[0].std().item() > 1e-3                       # a picture, not a constant
    # call 2, right after: fewer rays (the array still holds call 1 beyond them), ragged against the 128-point tiles
    o2, d2 = _inputs(301, 3)
    ref2, got2 = _render_both(L, P, st, packed, ladder, s, rays=(o2, d2))
    same(ref2, got2, "call 2")
    # call 3: camera mode (the generated rays go through the array too), a shard of a 40 x 40 frame
    pose = g1.pose_spherical(30.0, -30.0, 4.0311).cuda()
    cam = (pose[None, :3, 3].contiguous(), pose[None, :3, :3].contiguous(), 1, 40, 40, g1.TINY_NERF_FOCAL * 0.4)
    ref3, got3 = _render_both(L, P, st, packed, ladder, s, camera=cam, ray_begin=123, b=1000)
    same(ref3, got3, "call 3")
    # ... and call 1 again gives call 1's result again (nothing of calls 2-3 leaks into it)
    # The outputs exist before the driver's free-memory count is read, and nothing but the entry point runs between the two
    # readings: a cudaMalloc inside it (the 2.7 MB workspace of this size, say) would show.
    px = torch.empty(2048, 3, device="cuda"); px2 = torch.empty(301, 3, device="cuda")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    _lib.check(L.nerf_g1_render_fused_static(P(packed), P(o1), P(d1), None, None, 0, 0, 0, 1.0, 0, 2048, s, 2.0, 6.0, P(ladder), P(u1), 0, 0, 0,
                                             0.0, 6.0, 1, P(px), None, None, st()), "render (static), repeat")
    _lib.check(L.nerf_g1_render_fused_static(P(packed), P(o2), P(d2), None, None, 0, 0, 0, 1.0, 0, 301, s, 2.0, 6.0, P(ladder), None, 0, 0, 0,
                                             0.0, 6.0, 1, P(px2), None, None, st()), "render (static), repeat 2")
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    assert torch.equal(px, got1[0]) and torch.equal(px2, got2[0])
    assert free_before - free_after <= (1 << 20), (free_before, free_after)

    # beyond the limit: a clear refusal, no launch
    big = torch.zeros(2049, 3, device="cuda")
    rc = L.nerf_g1_render_fused_static(P(packed), P(big), P(big), None, None, 0, 0, 0, 1.0, 0, 2049, s, 2.0, 6.0, P(ladder), None, 0, 0, 0,
                                       0.0, 6.0, 1, P(px), None, None, st())
    assert rc == -3 and b"131072" in L.nerf_last_error()
    with pytest.raises(ValueError):
        _lib.check(rc, "nerf_g1_render_fused_static")
    assert L.nerf_g1_render_fused_static(P(packed), P(big), P(big), None, None, 0, 0, 0, 1.0, 0, 64, 16, 2.0, 6.0, None, None, 0, 0, 0,
                                         0.0, 6.0, 1, P(px), None, None, st()) == -2          # S < 32: as the existing entry point
    assert np.isfinite(px.cpu().numpy()).all()
