"""CPU-side checks: the C-ABI library loads and exports every symbol of include/nerf_b200.h,
host logic (flat parameters, layouts, dataset index decode, ray sharding) and the world_size-2
gloo path.  No kernel is launched here."""
import os
import re
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def test_library_exports_every_declared_symbol():
    import ctypes
    from nerf_b200 import _lib
    if not os.path.exists(_lib.library_path()):
        from nerf_b200 import build
        build.build()
    header = open(os.path.join(ROOT, "include", "nerf_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(nerf_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 20
    handle = ctypes.CDLL(_lib.library_path())
    for name in declared:
        assert hasattr(handle, name), f"{name} declared in the header but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert _lib.lib().nerf_abi_version() == 1


def test_param_layout_c_vs_python():
    from nerf_b200 import ops
    from nerf_b200.model import g1_param_layout_py
    for dx, dd in [(3, 3), (7, 3), (5, 4)]:
        assert ops.g1_param_layout(dx, dd) == g1_param_layout_py(dx, dd)
    assert ops.g1_param_layout(3, 3)[1] == 638468


def test_module_flat_storage_and_checkpoint(nerf_pth_state_dict):
    from nerf_b200.model import NeRF
    m = NeRF(normalize_position=6.0)
    assert list(m.state_dict().keys()) == list(nerf_pth_state_dict.keys())
    m.load_state_dict(nerf_pth_state_dict, strict=True)
    assert m._flat_ok()
    flat = m.flat_parameters()
    off = m._offsets
    for (k, v), o in zip(nerf_pth_state_dict.items(), off):
        assert torch.equal(flat[o:o + v.numel()].view(v.shape), v)
    # an in-place optimiser update through the parameters is visible in the flat vector
    with torch.no_grad():
        m.color.bias.add_(1.0)
    assert torch.equal(flat[off[3]:off[3] + 3], nerf_pth_state_dict["color.bias"] + 1.0)
    # round trip through torch.save / load (the reference's checkpoint path, nb cell 9)
    import io
    buf = io.BytesIO(); torch.save(m.state_dict(), buf); buf.seek(0)
    m2 = NeRF(normalize_position=6.0); m2.load_state_dict(torch.load(buf), strict=True)
    assert torch.equal(m2.flat_parameters(), flat)
    m3 = m.double().float()          # _apply re-flattens
    assert m3._flat_ok()


def test_no_cpu_fallback():
    from nerf_b200 import _lib
    from nerf_b200.model import NeRF
    m = NeRF(normalize_position=6.0)
    with pytest.raises(_lib.NerfB200Error):
        m.render_rays(torch.zeros(4, 3), torch.ones(4, 3), 2.0, 6.0, 8)


def test_api_signatures():
    import inspect
    from nerf.model import NeRF
    from nerf.dataset import PixelRayDataset  # noqa: F401
    sig = inspect.signature(NeRF.render_rays)
    assert list(sig.parameters)[:6] == ["self", "rays_o", "rays_d", "near", "far", "num_samples"]
    for kw in ("states_x", "states_d", "randomly_sample", "density_noise_std"):
        assert kw in sig.parameters
    sig = inspect.signature(NeRF.render_image)
    assert list(sig.parameters)[:9] == ["self", "camera_o", "camera_r", "image_h", "image_w", "focal_length",
                                        "near", "far", "num_samples"]
    for name in ("generate_rays", "rays_to_world_coordinates", "alpha_compositing_coefficients",
                 "spherical_to_cartesian", "get_rotation_matrix", "direction_to_rotation_matrix"):
        assert hasattr(NeRF, name)
    assert NeRF(normalize_position=6.0).stages_axis is False           # G1: [B,3]
    assert NeRF(normalize_position=20.0, density_inputs=7).stages_axis  # G2: [B,stages,3]


def test_camera_helpers_vs_reference_formulae():
    from nerf.model import NeRF
    yaw, el = torch.tensor([0.3, 1.1]), torch.tensor([-0.2, 0.7])
    v = NeRF.spherical_to_cartesian(yaw, el)
    np.testing.assert_allclose(v.norm(dim=-1).numpy(), 1.0, atol=1e-6)
    r = NeRF.direction_to_rotation_matrix(torch.tensor([[0.0, 0.0, -1.0], [1.0, 2.0, -0.5]]))
    for m in r:
        np.testing.assert_allclose((m.T @ m).numpy(), np.eye(3), atol=1e-5)
    # camera -z (the viewing axis of generate_rays) maps to the requested direction
    d = torch.tensor([1.0, 2.0, -0.5]); d = d / d.norm()
    np.testing.assert_allclose((r[1] @ torch.tensor([0.0, 0.0, -1.0])).numpy(), d.numpy(), atol=1e-5)


def test_dataset_index_decode_and_items():
    from hypothesis import given, settings, strategies as st
    from nerf.dataset import PixelRayDataset
    from oracle import g1

    @settings(max_examples=200, deadline=None)
    @given(st.integers(1, 9), st.integers(1, 17), st.integers(1, 13), st.data())
    def prop(n, h, w, data):
        idx = data.draw(st.integers(0, n * h * w - 1))
        bi, hi, wi = PixelRayDataset.decode_index(idx, n, h, w)
        assert (bi * h + hi) * w + wi == idx and 0 <= bi < n and 0 <= hi < h and 0 <= wi < w
    prop()
    g = torch.Generator().manual_seed(0)
    images, poses = torch.rand(3, 6, 5, 3, generator=g), torch.randn(3, 4, 4, generator=g)
    ds = PixelRayDataset(images, poses, 4.4)
    assert len(ds) == 90
    it = ds[57]
    bi, hi, wi = 57 // 30, (57 // 5) % 6, 57 % 5
    assert torch.equal(it["pixels"], images[bi, hi, wi]) and torch.equal(it["rays_o"], poses[bi, :3, 3])
    assert torch.equal(it["pose_d"], it["rays_d"])                       # the reference's quirk, dataset.py:315
    rays = g1.generate_rays(6, 5, 4.4)
    _, d = g1.rays_to_world_coordinates(rays[hi, wi], poses[bi, :3, 3], poses[bi, :3, :3])
    assert torch.equal(it["rays_d"], d)
    # HEAD constructor form with a segmentation tensor
    ds2 = PixelRayDataset(images, torch.zeros(3, 6, 5, dtype=torch.long), poses, 4.4)
    assert "label" in ds2[3] and ds2.focal_length == 4.4
    # G2 keyword form
    ds3 = PixelRayDataset(images, poses[:, :3], 4.4, states_x=torch.rand(3, 2))
    assert ds3[0]["states_x"].shape == (2,)


def test_shard_range_partitions():
    from nerf_b200.parallel import shard_range
    for total in (0, 1, 7, 640000, 10001):
        for world in (1, 2, 3, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c and a <= b
            assert sum(b - a for a, b in spans) == total


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    from nerf_b200.model import NeRF
    from nerf_b200.parallel import broadcast_parameters, shard_range
    torch.manual_seed(100 + rank)                       # different init on every rank
    m = NeRF(normalize_position=6.0)
    broadcast_parameters(m, src=0)
    flat = m.flat_parameters()
    ref = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(ref, flat)
    same = all(torch.equal(ref[0], r) for r in ref)
    # the Trainer's data-parallel rule: sum of per-rank mean-loss grads * (1/world) == grad of the global mean
    g = torch.Generator().manual_seed(7)
    x = torch.randn(8, 5, generator=g); w = torch.randn(5, generator=g)
    lo, hi = shard_range(8, rank, world)
    wl = w.clone().requires_grad_(True)
    ((x[lo:hi] @ wl) ** 2).mean().backward()
    grad = wl.grad.clone(); dist.all_reduce(grad); grad *= 1.0 / world
    wg = w.clone().requires_grad_(True); ((x @ wg) ** 2).mean().backward()
    # dataset shards: disjoint, complete
    from nerf_b200.dataset import PixelRayDataset
    ds = PixelRayDataset(torch.zeros(2, 4, 3, 3), torch.zeros(2, 4, 4), 1.0)
    order = torch.randperm(len(ds), generator=torch.Generator().manual_seed(0))
    mine = order[rank::world]
    allidx = [torch.empty(len(ds) // world, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(allidx, mine)
    cover = torch.cat(allidx).sort()[0]
    q.put((rank, same, torch.allclose(grad, wg.grad, atol=1e-6), torch.equal(cover, torch.arange(len(ds)))))
    dist.destroy_process_group()


def test_gloo_world_size_2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert len(res) == 2 and all(all(r[1:]) for r in res), res


def _image_ray_inputs():
    g = torch.Generator().manual_seed(11)            # same construction as tests/golden/make_image_ray_dataset.py
    images = torch.rand(3, 16, 24, 3, generator=g)
    poses = torch.eye(4).repeat(3, 1, 1); poses[:, :3, :] = torch.randn(3, 3, 4, generator=g)
    states = torch.randn(3, 5, generator=g)
    return images, poses, states


def test_image_ray_dataset_matches_reference_items():
    """Block-stratified items against the UNMODIFIED reference class (fixture generated from it):
    same length, same keys, same indices (global-generator draw included), same tensors."""
    from nerf.dataset import ImageRayDataset
    z = np.load(os.path.join(GOLDEN, "image_ray_dataset.npz"))
    images, poses, states = _image_ray_inputs()
    ds = ImageRayDataset(images, poses, states, 20.0, num_vertical_blocks=4, num_horizontal_blocks=4, num_samples_per_block=2)
    assert len(ds) == int(z["len"])
    for idx in (0, 17, 35):
        torch.manual_seed(100 + idx)
        item = ds[idx]
        keys = sorted(k.split("/", 1)[1] for k in z.files if k.startswith(f"{idx}/"))
        assert sorted(item.keys()) == keys
        for k in keys:
            want = z[f"{idx}/{k}"]
            assert tuple(item[k].shape) == want.shape, k
            if want.dtype.kind == "i":
                assert np.array_equal(item[k].numpy(), want), k
            else:
                np.testing.assert_allclose(item[k].numpy(), want, rtol=0, atol=1e-6, err_msg=k)


def test_render_image_launch_pieces_cover_and_keep_the_kernel_choice():
    """NeRF._launch_pieces: a shard's launches cover it exactly, in order; outside the frame's short tail chunk every launch
    has at least COMPOSITE_WARP_BELOW rays (the thread-per-ray composite, as in the frame) and stays outside the tail;
    the tail's launches stay below it (the warp-per-ray composite, as in the frame)."""
    from nerf_b200 import ops
    from nerf_b200.model import NeRF
    thr = ops.COMPOSITE_WARP_BELOW
    for total, chunk in ((640000, 32768), (640000, 1 << 20), (40000, 32768), (10000, 32768), (10000, 1024), (70000, 32768),
                         (65536, 32768), (82020, 32768), (16384, 16384), (16385, 16384), (3 * 32768 + 5, 32768), (2 * 640000, 1 << 20)):
        last = total % chunk
        tail0 = total - last if (chunk >= thr and 0 < last < thr) else total
        for world in (1, 2, 3, 8, 64):
            for r in range(world):
                rb, re = total * r // world, total * (r + 1) // world
                x = rb
                for lo, hi, i, j in NeRF._launch_pieces(rb, re, total, chunk):
                    assert 0 <= lo <= i < j <= hi <= total and i == x
                    x = j
                    if chunk < thr:
                        assert hi - lo <= chunk
                    elif i < tail0:
                        assert hi - lo >= thr and hi <= tail0
                    else:
                        assert lo >= tail0 and hi - lo < thr
                assert x == re
    assert list(NeRF._launch_pieces(5, 5, 100, 32768)) == []                     # an empty shard
