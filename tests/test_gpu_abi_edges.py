"""Error behaviour and degenerate sizes of the C ABI (include/nerf_b200.h): status codes instead of
crashes, zero-ray calls are no-ops, undersized workspaces and unsupported shapes are refused."""
import pytest
import torch

pytestmark = pytest.mark.gpu

BAD_ARG, UNSUPPORTED, WORKSPACE = -1, -2, -3


@pytest.fixture(scope="module")
def L():
    from nerf_b200 import _lib
    return _lib.lib()


@pytest.fixture(scope="module")
def packed(nerf_pth_state_dict):
    from nerf_b200 import ops
    from nerf_b200.model import NeRF
    m = NeRF(normalize_position=6.0, precision="bf16").cuda(); m.load_state_dict(nerf_pth_state_dict)
    return ops.g1_pack_weights(m.flat_parameters())


def _render_args(P, packed, o, d, b, s, ws, nbytes, pixels):
    from nerf_b200 import _lib
    return (P(packed), 3, 3, P(o), P(d), None, None, 0, 0, 0, 1.0, 0, b, s, 2.0, 6.0, None, None, None, 0, 0, 0, None, 0.0,
            6.0, 1, None, 0, 1, P(pixels), None, None, None, None, None, None, P(ws), nbytes, _lib.stream())


def test_render_fused_conditional_status_codes(L):
    """Dx > 3 needs states_x with enough rows for every ray of the call; the workspace grows by the bias rows."""
    from nerf_b200 import _lib, ops
    from nerf_b200.model import NeRF
    P = _lib.ptr
    torch.manual_seed(0)
    m = NeRF(normalize_position=6.0, density_inputs=7, precision="bf16").cuda()
    pk = ops.g1_pack_weights(m.flat_parameters(), 7, 3)
    o = torch.randn(8, 3, device="cuda"); d = torch.randn(8, 3, device="cuda"); px = torch.empty(8, 3, device="cuda")
    sx = torch.rand(8, 4, device="cuda")
    n = L.nerf_g1_render_workspace_bytes(8, 64, 7, 8)
    assert n > L.nerf_g1_render_workspace_bytes(8, 64, 3, 0) > 0
    assert L.nerf_g1_render_workspace_bytes(8, 64, 2, 0) == BAD_ARG
    ws = torch.empty(n, dtype=torch.uint8, device="cuda")
    def call(states, n_states, per, nbytes=n):
        return L.nerf_g1_render_fused(P(pk), 7, 3, P(o), P(d), None, None, 0, 0, 0, 1.0, 0, 8, 64, 2.0, 6.0, None, None, None, 0, 0, 0,
                                      None, 0.0, 6.0, 1, P(states), n_states, per, P(px), None, None, None, None, None, None,
                                      P(ws), nbytes, _lib.stream())
    assert call(sx, 8, 1) == 0
    assert call(sx, 2, 4) == 0                                   # 4 rays per state row
    assert call(None, 8, 1) == BAD_ARG                           # conditional model without states
    assert call(sx, 4, 1) == BAD_ARG                             # rays 4..7 have no row
    assert call(sx, 8, 0) == BAD_ARG
    assert call(sx, 8, 1, L.nerf_g1_render_workspace_bytes(8, 64, 3, 0)) == WORKSPACE
    torch.cuda.synchronize()


def test_render_fused_status_codes(L, packed):
    from nerf_b200 import _lib
    P = _lib.ptr
    o = torch.randn(8, 3, device="cuda"); d = torch.randn(8, 3, device="cuda"); px = torch.empty(8, 3, device="cuda")
    n = L.nerf_g1_render_workspace_bytes(8, 64, 3, 0)
    ws = torch.empty(n, dtype=torch.uint8, device="cuda")
    assert L.nerf_g1_render_fused(*_render_args(P, packed, o, d, 8, 64, ws, n, px)) == 0
    assert L.nerf_g1_render_fused(*_render_args(P, packed, o, d, 0, 64, ws, n, px)) == 0            # zero rays: no-op
    assert L.nerf_g1_render_fused(*_render_args(P, packed, o, d, 8, 16, ws, n, px)) == UNSUPPORTED  # S < 32: fp32 path only
    assert L.nerf_g1_render_fused(*_render_args(P, packed, o, d, 8, 64, ws, n - 1, px)) == WORKSPACE
    assert L.nerf_g1_render_fused(*_render_args(P, None, o, d, 8, 64, ws, n, px)) == BAD_ARG
    a = list(_render_args(P, packed, o, d, 8, 64, ws, n, px)); a[4] = None                           # rays_o without rays_d
    assert L.nerf_g1_render_fused(*a) == BAD_ARG
    with pytest.raises(ValueError):
        _lib.check(UNSUPPORTED, "x")
    torch.cuda.synchronize()


def test_composite_zero_rays_and_bad_args(L):
    from nerf_b200 import _lib
    P = _lib.ptr
    s = torch.zeros(4, 64, device="cuda"); rgb = torch.zeros(4, 64, 3, device="cuda"); t = torch.zeros(4, 64, device="cuda")
    o = torch.zeros(4, 3, device="cuda"); px = torch.empty(4, 3, device="cuda")
    assert L.nerf_composite_fwd(P(s), P(rgb), None, P(o), P(o), P(t), 0, 64, P(px), None, None, None, _lib.stream()) == 0
    assert L.nerf_composite_fwd(None, P(rgb), None, P(o), P(o), P(t), 4, 64, P(px), None, None, None, _lib.stream()) == BAD_ARG
    assert L.nerf_composite_fwd(P(s), P(rgb), None, P(o), P(o), P(t), 4, 0, P(px), None, None, None, _lib.stream()) == BAD_ARG


def test_train_entry_status_codes(L, packed):
    from nerf_b200 import _lib
    P = _lib.ptr
    b, s = 4, 64
    o = torch.randn(b, 3, device="cuda"); e = lambda *sh: torch.empty(*sh, device="cuda")
    n = L.nerf_g1_train_workspace_bytes(b, s, 3)
    assert n > 0 and L.nerf_g1_train_workspace_bytes(b, s, 2) == BAD_ARG and L.nerf_g1_train_workspace_bytes(0, s, 3) == BAD_ARG
    assert L.nerf_g1_train_workspace_bytes(b, s, 7) > n                                           # conditional inputs need more
    ws = torch.empty(n, dtype=torch.uint8, device="cuda")
    args = lambda dx, sx, nbytes: (P(packed), dx, 3, P(o), P(o), b, s, 2.0, 6.0, None, None, None, 0, 0, 0, 6.0, 1, sx,
                                   P(e(b, s)), P(e(b, s)), P(e(b, s, 3)), P(ws), nbytes, _lib.stream())
    assert L.nerf_g1_train_fwd_tc(*args(3, None, n)) == 0
    assert L.nerf_g1_train_fwd_tc(*args(3, None, n - 1)) == WORKSPACE
    assert L.nerf_g1_train_fwd_tc(*args(7, None, n)) == BAD_ARG                                   # states missing
    torch.cuda.synchronize()


def test_python_layer_refuses_cpu_tensors_and_bad_shapes():
    from nerf_b200 import _lib, ops
    with pytest.raises(_lib.NerfB200Error):
        ops.composite(torch.zeros(2, 8), torch.zeros(2, 8, 3), rays_o=torch.zeros(2, 3), rays_d=torch.zeros(2, 3), t=torch.zeros(2, 8))
    from nerf_b200.head import HeadNeRF
    with pytest.raises(ValueError):
        HeadNeRF(hidden_size=128)
    with pytest.raises(ValueError):
        HeadNeRF(encoding_size=16, precision="bf16")


def test_zero_ray_render_through_the_model(nerf_pth_state_dict):
    from nerf_b200.model import NeRF
    for prec in ("fp32", "bf16"):
        m = NeRF(normalize_position=6.0, precision=prec).cuda(); m.load_state_dict(nerf_pth_state_dict)
        with torch.no_grad():
            out = m.render_rays(torch.zeros(0, 3, device="cuda"), torch.zeros(0, 3, device="cuda"), 2.0, 6.0, 64)
        assert out.shape == (0, 3)
