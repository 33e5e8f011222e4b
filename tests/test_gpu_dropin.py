"""Drop-in proof on the reference's OWN entry points: the unmodified
`train_conditional_nerf.py` (G2 API) and the unmodified code cells of `examples/example.ipynb` (G1 API)
run against this repo's `nerf/` package, and their `render_rays -> loss.backward() -> optim.Adam.step()`
loop reaches the tcgen05 kernels (`ops._G1TrainTC`) when NERF_B200_PRECISION=bf16.

The entry points are read from /root/reference when it exists, else from baseline/_ref/ (byte copies staged
by scripts/stage_reference.py at build time; that directory is git-ignored and travels to the GPU box).
What is executed is pinned by SHA-256 below, so "unmodified" is a checked fact, not a claim."""
import hashlib
import json
import os
import pickle
import subprocess
import sys
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TRAIN_SHA = "62b44d796f9f374aee404c8b3956406d5f1257c3e9f0243ab73a4c53eb64eb87"     # /root/reference/train_conditional_nerf.py
CELL_SHA = {0: "d53f6d55e1a6eb12cc06385c42627ddf14ecde3e7247cbad3730b211c87eb15a", 1: "acbb03bb9779959bfbb04921b053315185c7e52785aaca1308f65d27e13a14d7",
            2: "43f110d4935faa657ad64f14696945c5d00d43772469d4ad47ee077e93fd59fc", 3: "71f253593ac920addfe480eafc2d70c3b3962083ccc6bdd441f59076037e90fb",
            4: "871aab90687dfafced5e26e1b20605b10eaa1eb6af18efd4e6304e043ec0401f", 5: "cf73e42fac9d99719b02228d76b747eb4ae541aef6ed1617bdacb3c6db7c69c7",
            6: "afd7787b3a224feb10ba3a693ec9771f564cfca53594e13e0233247ef8ccab57", 7: "1d308263b5026465cd6703bad2728a82c1fd55025df78d97c294fe8adb4cfb13",
            8: "6afa5455eb5268bc911fa6bb43c7f3823cc17581cc31f51538c97a1bb33edc39", 9: "2c39f8e9be8e398ce63ce2495bcb4bf18b21ca0ffd37da1ecea5b838eedc0bb8"}


def _reference_train_script():
    for p in ("/root/reference/train_conditional_nerf.py", os.path.join(ROOT, "baseline", "_ref", "train_conditional_nerf.py")):
        if os.path.exists(p):
            return p
    pytest.skip("reference entry point not staged (scripts/stage_reference.py runs in __graft_entry__.build())")


def _reference_cells():
    p = os.path.join(ROOT, "baseline", "_ref", "example_cells.json")
    if os.path.exists("/root/reference/examples/example.ipynb"):
        nb = json.load(open("/root/reference/examples/example.ipynb"))
        return ["".join(c["source"]) for c in nb["cells"] if c["cell_type"] == "code"]
    if os.path.exists(p):
        return json.load(open(p))
    pytest.skip("reference notebook cells not staged (scripts/stage_reference.py)")


def _orbit(n, radius=4.0311, seed=0):
    g = np.random.default_rng(seed)
    th = g.uniform(0, 2 * np.pi, n); ph = np.deg2rad(g.uniform(20, 50, n))
    return np.stack([radius * np.cos(th) * np.cos(ph), radius * np.sin(th) * np.cos(ph), radius * np.sin(ph)], -1).astype(np.float32)


def test_unmodified_reference_train_script_reaches_tcgen05(tmp_path, nerf_pth_state_dict):
    script = _reference_train_script()
    assert hashlib.sha256(open(script, "rb").read()).hexdigest() == TRAIN_SHA, "the staged entry point is not the reference's file"
    from nerf_b200.model import NeRF as G1
    n, h, w, sx = 5, 16, 16, 4
    pos = _orbit(n)
    look = -pos / np.linalg.norm(pos, axis=-1, keepdims=True)
    states = np.random.default_rng(1).uniform(-3.0, 3.0, (n, sx)).astype(np.float32)        # the script divides by pi
    # data_for_nerf.pkl is a missing blob: synthesise it with the reference's keys.  A multi-view consistent
    # scene: the views of examples/nerf.pth from the script's own cameras (direction_to_rotation_matrix, focal = W 50/36)
    teacher = G1(normalize_position=6.0, precision="fp32").cuda(); teacher.load_state_dict(nerf_pth_state_dict)
    with torch.no_grad():
        cam_r = G1.direction_to_rotation_matrix(torch.from_numpy(look).cuda())
        images = teacher.render_image(torch.from_numpy(pos).cuda(), cam_r, h, w, w * 50.0 / 36.0, 2.0, 6.0, 64).cpu().numpy()
    assert images.std() > 0.05                                      # the cameras see the object
    data = dict(images=images, poses=np.concatenate([pos, look], -1), states=states)
    pkl_path = tmp_path / "data_for_nerf.pkl"
    with open(pkl_path, "wb") as f:
        pickle.dump(data, f)
    stats = tmp_path / "stats.json"
    logdir = tmp_path / "experiment"
    epochs, batch = 40, 256                                         # (n-1)*h*w = 1024 pixels -> 4 batches per epoch, 160 iterations
    losses = tmp_path / "losses.json"
    # The script is executed as it is (runpy); what surrounds it only seeds torch (the script never does) and records the value of
    # every scalar `.backward()` is called on -- the script computes its loss and drops it.
    driver = (
        "import sys, runpy, json, atexit, torch\n"
        "torch.manual_seed(0)\n"
        "_l, _bw = [], torch.Tensor.backward\n"
        "def _rec(self, *a, **k):\n"
        "    if self.numel() == 1: _l.append(float(self.detach()))\n"
        "    return _bw(self, *a, **k)\n"
        "torch.Tensor.backward = _rec\n"
        "atexit.register(lambda: json.dump(_l, open(%r, 'w')))\n"
        "sys.path.insert(0, %r); sys.argv = %r\n"
        "runpy.run_path(%r, run_name='__main__')\n" % (
            str(losses), ROOT, [script, "--logging-dir", str(logdir), "--data", str(pkl_path), "--epochs", str(epochs), "--batch-size", str(batch),
                                "--log-interval", "50", "--near-plane", "2.0", "--far-plane", "6.0", "--normalize-position", "6.0",
                                "--learning-rate", "0.0005", "--num-samples-per-ray", "64"], script))
    env = dict(os.environ, NERF_B200_PRECISION="bf16", NERF_B200_STATS=str(stats), PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, "-c", driver], env=env, capture_output=True, text=True, timeout=900, cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    c = json.load(open(stats))
    iters = epochs * 4
    assert c.get("g1_train_fwd_tc") == iters and c.get("g1_train_bwd_tc") == iters, c     # every loss.backward() ran the tcgen05 backward
    assert c.get("g1_mlp_fwd_fp32", 0) == 0, c                                           # and nothing fell to the FFMA path
    assert c.get("g1_render_fused", 0) >= 4, c                                           # the no_grad validation renders
    psnrs = np.load(logdir / "psnrs.npy"); its = np.load(logdir / "iternums.npy")
    assert list(its) == [0, 50, 100, 150] and np.isfinite(psnrs).all()
    ls = np.array(json.load(open(losses)))
    first, last = ls[:20].mean(), ls[-20:].mean()
    print(f"reference train script on tcgen05: loss {first:.4f} (first 20 steps) -> {last:.4f} (last 20 of {len(ls)}); "
          f"held-out PSNR at iterations {list(map(int, its))} = {[round(float(p), 2) for p in psnrs]}")
    assert len(ls) == iters and np.isfinite(ls).all()
    assert last < 0.75 * first, (first, last)                                            # the loss falls
    sd = torch.load(logdir / "model.pth", map_location="cpu")
    assert len(sd) == 44 and sd["block_0.0.weight"].shape == (256, 20 * (3 + sx))
    imgs = np.load(logdir / "rendered_images.npy")
    assert imgs.shape == (4, 1, h, w, 3) and json.load(open(logdir / "params.json"))["batch_size"] == batch


@pytest.mark.parametrize("precision", ["bf16", "bf16x3"])
def test_unmodified_notebook_cells_reach_tcgen05(tmp_path, monkeypatch, nerf_pth_state_dict, precision):
    """example.ipynb cells 0-9 as they are, on the bf16 tensor-core step and on the fp32-grade one (bf16x3)"""
    cells = _reference_cells()
    fam = "tc" if precision == "bf16" else "x3"
    for i, sha in CELL_SHA.items():
        assert hashlib.sha256(cells[i].encode()).hexdigest() == sha, f"cell {i} is not the reference's"
    from nerf_b200 import ops
    from nerf_b200.model import NeRF as G1
    from oracle import g1 as og
    # tiny_nerf_data.npz is a missing blob of the reference: synthesise it with the reference's keys;
    # views of examples/nerf.pth rendered by the fp32 path at 12 x 12
    n, h, w = 5, 12, 12
    focal = og.TINY_NERF_FOCAL * w / 100
    teacher = G1(normalize_position=6.0, precision="fp32").cuda(); teacher.load_state_dict(nerf_pth_state_dict)
    poses = torch.stack([og.pose_spherical(40.0 * i, -30.0, 4.0311) for i in range(n)])
    with torch.no_grad():
        images = teacher.render_image(poses[:, :3, 3].cuda(), poses[:, :3, :3].cuda(), h, w, focal, 2.0, 6.0, 64).cpu().numpy()
    np.savez(tmp_path / "tiny_nerf_data.npz", images=images, poses=poses.numpy(), focal=np.float64(focal))
    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv("NERF_B200_PRECISION", precision)
    plt = types.ModuleType("matplotlib.pyplot")                     # matplotlib is not installed: the plotting calls become no-ops
    for name in ("figure", "subplot", "title", "imshow", "plot", "show", "close"):
        setattr(plt, name, lambda *a, **k: None)
    mpl = types.ModuleType("matplotlib"); mpl.pyplot = plt
    monkeypatch.setitem(sys.modules, "matplotlib", mpl); monkeypatch.setitem(sys.modules, "matplotlib.pyplot", plt)
    before = dict(ops.COUNTERS)
    torch.manual_seed(0)
    ns = {"__name__": "__notebook__"}
    for i in range(10):                                             # cells 0-9: imports, data, dataset, model, optimiser, the loop, the save
        exec(compile(cells[i], f"<example.ipynb cell {i}>", "exec"), ns)
    fwd = ops.COUNTERS[f"g1_train_fwd_{fam}"] - before.get(f"g1_train_fwd_{fam}", 0)
    bwd = ops.COUNTERS[f"g1_train_bwd_{fam}"] - before.get(f"g1_train_bwd_{fam}", 0)
    assert fwd == bwd == 100, (fwd, bwd)                            # 100 epochs x 1 batch of (n-1)*144 = 576 pixels
    assert ops.COUNTERS["g1_mlp_fwd_fp32"] == before.get("g1_mlp_fwd_fp32", 0)     # nothing fell to the FFMA path
    nerf = ns["nerf"]
    assert nerf.precision == precision and ns["iteration"] == 99
    with torch.no_grad():
        out = nerf.render_image(ns["test_pose"][..., :3, 3], ns["test_pose"][..., :3, :3], h, w, focal, 2.0, 6.0, 64)
        psnr_end = (-10.0 * torch.log(((out - ns["test_image"]) ** 2).mean()) / 2.30258509299).item()
    psnr_0 = float(ns["psnrs"][0])
    print(f"notebook loop on tcgen05 ({precision}): held-out PSNR {psnr_0:.2f} dB after 1 step -> {psnr_end:.2f} dB after 100")
    assert psnr_end > psnr_0 + 0.3
    sd = torch.load(tmp_path / "nerf.pth", map_location="cpu")      # cell 9
    assert list(sd.keys()) == list(nerf_pth_state_dict.keys())


def test_reference_loop_inline_matches_trainer_gradients(nerf_pth_state_dict):
    """The same node, checked numerically: gradients of `((render_rays(...) - target)**2).mean().backward()`
    on the bf16 autograd path equal the fused Trainer's flat gradient for the same draws (same kernels, the
    loss/composite backward taken by different launches), and torch.optim.Adam moves the flat parameters."""
    from nerf_b200.model import NeRF as G1
    from nerf_b200.train import Trainer
    from oracle import g1 as og
    torch.manual_seed(0)
    m = G1(normalize_position=6.0, precision="bf16").cuda(); m.load_state_dict(nerf_pth_state_dict)
    b, s = 300, 64
    pose = og.pose_spherical(30.0, -30.0, 4.0311)
    rays = og.generate_rays(20, 15, 40.0).reshape(-1, 3)
    o, d = og.rays_to_world_coordinates(rays, pose[:3, 3].expand_as(rays), pose[:3, :3].expand(b, 3, 3))
    o, d = o.cuda().contiguous(), d.cuda().contiguous()
    target = torch.rand(b, 3, device="cuda")
    u = torch.rand(b, s, device="cuda"); noise = torch.randn(b, s, 1, device="cuda")
    out = m.render_rays_full(o, d, 2.0, 6.0, s, u=u, noise=noise)
    loss = ((out["pixels"] - target) ** 2).mean()
    m.zero_grad(); loss.backward()
    g_api = torch.cat([p.grad.reshape(-1) for p in m.parameters()])
    tr = Trainer(m, precision="bf16")
    l2 = tr.step(o, d, target, u=u, noise=noise, apply_update=False)
    assert abs(loss.item() - l2.item()) <= 1e-6 * max(1.0, abs(l2.item()))
    rel = ((g_api - tr.flat_grad).norm() / tr.flat_grad.norm()).item()
    print(f"autograd-node gradient vs Trainer flat gradient: rel-L2 {rel:.2e}")
    assert rel <= 1e-5
    # public signature, G1 shape, RNG order of the reference (rand then randn), and an optimiser step through views of the flat buffer
    opt = torch.optim.Adam(m.parameters(), lr=1e-4)
    flat0 = m.flat_parameters().clone()
    px = m.render_rays(o, d, 2.0, 6.0, s, randomly_sample=True, density_noise_std=1.0)
    assert px.shape == (b, 3) and px.requires_grad
    opt.zero_grad(); ((px - target) ** 2).mean().backward(); opt.step()
    assert m._flat_ok() and (m.flat_parameters() - flat0).abs().max().item() > 0
