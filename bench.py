#!/usr/bin/env python
"""Headline benchmark: rays/s of the NeRF training step (fwd+bwd+Adam) on 4096-ray x 64-sample
batches (BASELINE.json configs[1]), plus render throughput, rooflines and the
reference's CPU path timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision fp32|bf16]

N > 1 is launched by torchrun (one rank per GPU, NCCL); weak scaling: 4096 rays per GPU.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

RAYS, SAMPLES, NEAR, FAR = 4096, 64, 2.0, 6.0
FLOP_PER_POINT_FWD = 1_261_568           # G1 MLP, unpadded MACs x 2
FLOP_PER_POINT_FWD_BWD = 3_704_832
COMPOSITE_FWD_BYTES_PER_RAY = 20 * SAMPLES + 20 + 24
COMPOSITE_BWD_BYTES_PER_RAY = 36 * SAMPLES + 40 + 24


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tensor_burst=d["bf16_tflops"], tensor_sustained=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tensor_burst=1590.0, tensor_sustained=1400.0, src="fallback")


def ncu_traffic():
    """dram__bytes_read + dram__bytes_write per launch of the training kernels from the committed ncu --set full
    summary (profiles/, written by scripts/ncu_summary.py at this workload size); {} when absent."""
    path = os.path.join(ROOT, "profiles", "r01_train_tc_v10_ncu.json")
    out = {}
    if os.path.exists(path):
        for k in json.load(open(path))["kernels"]:
            for name in ("g1_fused_tc_kernel", "g1_bwd_tc_kernel", "g1_wgrad_tc_kernel"):
                if name in k["kernel"] and "dram_bytes" in k:
                    out[name] = k["dram_bytes"]
    return out


def nerf_pth_state_dict():
    z = np.load(os.path.join(ROOT, "tests", "golden", "nerf_pth.npz"))
    return {str(k): torch.from_numpy(z[str(k)]) for k in z["__keys__"]}


def synthetic_batch(n_rays, seed):
    """D2: rays from orbit poses (theta uniform, phi in [-80,-10], r=4.0311) x 100x100 pixels."""
    from oracle.g1 import TINY_NERF_FOCAL
    g = torch.Generator().manual_seed(seed)
    th = torch.rand(n_rays, generator=g) * 2 * np.pi
    ph = (-80 + 70 * torch.rand(n_rays, generator=g)) * np.pi / 180
    px = torch.randint(0, 100, (n_rays, 2), generator=g).float()
    r = 4.0311
    # c2w = flip . rot_theta . rot_phi . trans_z(r), vectorised (oracle.g1.pose_spherical)
    ct, st, cp, sp = th.cos(), th.sin(), ph.cos(), ph.sin()
    z = torch.zeros_like(ct); o1 = torch.ones_like(ct)
    rot_phi = torch.stack([o1, z, z, z, cp, -sp, z, sp, cp], -1).view(-1, 3, 3)
    rot_th = torch.stack([ct, z, -st, z, o1, z, st, z, ct], -1).view(-1, 3, 3)
    flip = torch.tensor([[-1.0, 0, 0], [0, 0, 1], [0, 1, 0]])
    rot = flip @ rot_th @ rot_phi
    origin = rot @ torch.tensor([0.0, 0.0, r])
    cam = torch.stack([(px[:, 1] - 49.5) / TINY_NERF_FOCAL, -(px[:, 0] - 49.5) / TINY_NERF_FOCAL, -o1], -1)
    d = (rot * cam[:, None, :]).sum(-1)
    return origin.contiguous(), d.contiguous()


class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True); self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except (ValueError, IndexError):
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------- CPU baseline -----
def cpu_train_step_rays_per_s(n_rays, steps, warmup=1):
    """The reference's CPU path for this workload: the G1 oracle (eager PyTorch fp32,
    all host threads) doing the same train step (fwd + bwd + Adam) on `n_rays` rays."""
    from oracle import g1
    torch.set_num_threads(os.cpu_count())
    m = g1.G1NeRF(normalize_position=6.0); m.load_state_dict(nerf_pth_state_dict())
    opt = torch.optim.Adam(m.parameters(), lr=1e-4)
    o, d = synthetic_batch(n_rays, seed=0)
    target = torch.rand(n_rays, 3, generator=torch.Generator().manual_seed(3))
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        torch.manual_seed(i)
        px = m.render_rays(o, d, NEAR, FAR, SAMPLES, randomly_sample=True, density_noise_std=1.0)
        opt.zero_grad()
        ((px - target) ** 2).mean().backward()
        opt.step()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    return n_rays / float(np.median(times)), float(np.sum(times))


def run_reference(args):
    """--impl reference: the reference's own CPU implementation (oracle port; the G1
    source is not in the reference tree) on a bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample = 512
    t0 = time.perf_counter()
    rps, _ = cpu_train_step_rays_per_s(sample, args.steps, args.warmup)
    ms = 1e3 * sample / rps
    line = {"impl": "reference", "metric": "rays/sec (train fwd+bwd)", "value": rps, "unit": "rays/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "tiny NeRF train step fwd+bwd+Adam, 4096-ray batches x 64 samples (G1, nerf.pth)",
                       "sample": f"{sample} rays x {SAMPLES} samples per step"},
            "cpu_baseline": {"value": rps, "unit": "rays/s", "cores": os.cpu_count(), "kind": "port",
                             "sample": f"{sample} rays x {SAMPLES} samples per step, median of {args.steps}"},
            "e2e": {"value": rps, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": time.perf_counter() - t0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ our arm -----
def dominant_roofline(kernels, peaks):
    """The contract's `roofline` object for the longest kernel of the step, against the resource that binds it."""
    name, k = max(kernels.items(), key=lambda kv: kv[1]["ms"])
    r = k[k["bound"]]
    return {"bound": k["bound"], "kernel": name, "achieved": r["achieved"], "unit": r["unit"], "frac": r["frac"],
            "peak": peaks["hbm"] if k["bound"] == "hbm" else peaks["tensor_sustained"],
            "peak_source": peaks["src"] + (" HBM copy" if k["bound"] == "hbm" else " bf16 sustained"),
            "traffic": k["traffic"], "ms": k["ms"], "algorithmic_bytes": k["hbm"]["bytes"]}


def timed_steps(fn, steps, flush):
    """Each step bracketed by its own CUDA events on the current stream; L2 is flushed
    (a 256 MiB write) between steps, outside the timed brackets."""
    evs = []
    for i in range(steps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(i); b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    return [a.elapsed_time(b) for a, b in evs]


def run_ours(args):
    import torch.distributed as dist
    from nerf_b200 import _lib, ops
    from nerf_b200.model import NeRF
    from nerf_b200.train import Trainer
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: nerf_b200 has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    _lib.lib()
    peaks = load_peaks()
    dev = torch.device("cuda", local)

    model = NeRF(normalize_position=6.0, precision=args.precision).to(dev)
    model.load_state_dict(nerf_pth_state_dict())
    trainer = Trainer(model, lr=1e-4, near=NEAR, far=FAR, num_samples=SAMPLES, density_noise_std=1.0,
                      randomly_sample=True, seed=1234)
    # a pool of distinct batches, resident in HBM (value) and in pinned host memory (e2e)
    n_pool = 8
    pool_h = []
    for i in range(n_pool):
        o, d = synthetic_batch(RAYS, seed=1000 * rank + i)
        tgt = torch.rand(RAYS, 3, generator=torch.Generator().manual_seed(7 + i))
        pool_h.append(torch.stack([o, d, tgt]).pin_memory())      # one pinned [3, RAYS, 3] block per batch: one H2D copy
    pool_d = [tuple(b.to(dev)) for b in pool_h]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    K, W = args.steps, args.warmup

    def step_resident(i):
        o, d, tgt = pool_d[i % n_pool]
        return trainer.step(o, d, tgt)

    host_loss = torch.empty(1).pin_memory()

    def step_e2e(i):
        o, d, tgt = pool_h[i % n_pool].to(dev, non_blocking=True)   # rays_o, rays_d, target pixels: contiguous views
        loss = trainer.step(o, d, tgt)
        host_loss.copy_(loss, non_blocking=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(W):
        step_resident(i)
    barrier()
    with ClockSampler(local) as clk:                     # spans both timed regions (value and e2e)
        ms = timed_steps(step_resident, K, flush)
        barrier()
        total_ms = float(np.sum(ms))
        for i in range(max(W // 2, 1)):
            step_e2e(i)
        barrier()
        ms_e2e = timed_steps(step_e2e, K, flush)
        barrier()
        time.sleep(0.05)
    total_e2e = float(np.sum(ms_e2e))
    loss_val = float(host_loss.item())

    # ---- dominant kernel group timed alone: MLP fwd+bwd through the C ABI ----
    b, s = RAYS, SAMPLES
    o, d, tgt = pool_d[0]
    flat = model.flat_parameters()
    P = _lib.ptr
    L = _lib.lib()
    kernels = {}
    if args.precision == "bf16":
        z = trainer._tc_buffers(b, s, dev)
        ladder = model._ladder(NEAR, FAR, s, dev)
        grads = torch.zeros_like(flat)

        def k_fwd(i):
            _lib.check(L.nerf_g1_train_fwd_tc(P(z["packed"]), 3, 3, P(o), P(d), b, s, NEAR, FAR, P(ladder), None, None, 1, 1, 0,
                                              6.0, 1, None, P(z["t"]), P(z["sigma"]), P(z["rgb"]), P(z["ws"]), z["n"], _lib.stream()), "fwd")

        def k_bwd(stage):
            def f(i):
                _lib.check(L.nerf_g1_train_bwd_tc(P(z["packed"]), P(z["packed_bwd"]), P(flat), 3, 3, P(z["d_sigma"]), P(z["d_rgb"]),
                                                  b, s, P(grads), P(z["ws"]), z["n"], stage, _lib.stream()), "bwd")
            return f

        def mlp_only(i):
            k_fwd(i); k_bwd(0)(i)
        for i in range(3):
            mlp_only(i)
        n_it = min(max(K // 2, 5), 20)
        ms_fwd = float(np.mean(timed_steps(k_fwd, n_it, flush)))
        ms_dgrad = float(np.mean(timed_steps(k_bwd(1), n_it, flush)))
        ms_wg = float(np.mean(timed_steps(k_bwd(3), n_it, flush)))
        ms_mlp = float(np.mean(timed_steps(mlp_only, n_it, flush)))
        n_tiles = (b * s + 127) // 128
        wgrad_bytes = n_tiles * ((10 + 9 + 2) * 65536 + 3 * 16384 + 2 * 4096)     # dz + unit inputs + head inputs + encodings + dh
        act_bytes = n_tiles * 65536                                                # one unit's [128 x 256] bf16 slabs, all tiles
        fwd_bytes = 10 * act_bytes + 2 * n_tiles * 16384 + 10 * n_tiles * 128 * 8 + b * s * 20   # activations, encodings, stats, t/sigma/rgb
        dgrad_bytes = 20 * act_bytes + 10 * n_tiles * 128 * 8 + n_tiles * 4096 + b * s * 16      # n_l in, dz_l out, stats, dh image, d_sigma/d_rgb
        traffic = ncu_traffic()

        def kern(name, ms, flop, nbytes):
            tf, gbs = flop / ms / 1e9, nbytes / ms / 1e6
            ft, fh = tf / peaks["tensor_sustained"], gbs / peaks["hbm"]
            return {"ms": ms, "bound": "tensor" if ft >= fh else "hbm", "frac": max(ft, fh),
                    "tensor": {"achieved": tf, "unit": "TFLOP/s", "frac": ft}, "hbm": {"achieved": gbs, "unit": "GB/s", "frac": fh, "bytes": nbytes},
                    "traffic": traffic.get(name)}
        kernels = {
            "g1_fused_tc_kernel<2,train>": kern("g1_fused_tc_kernel", ms_fwd, FLOP_PER_POINT_FWD * b * s, fwd_bytes),
            "g1_bwd_tc_kernel<2>": kern("g1_bwd_tc_kernel", ms_dgrad, 2 * 9 * 65536 * b * s, dgrad_bytes),
            "g1_wgrad_tc_kernel": kern("g1_wgrad_tc_kernel", ms_wg, 2 * (9 * 65536 + 3 * 16384 + 10 * 256 + 2 * 1024) * b * s, wgrad_bytes),
        }
        mlp_kernel_name = "G1 MLP fwd+bwd on tcgen05 (g1_fused_tc_kernel<train> + g1_bwd_tc_kernel + g1_wgrad_tc_kernel + unfold)"
    else:
        t = ops.sample_linear(NEAR, FAR, s, b, dev, jitter=True, seed=1)
        ex, ed, _ = ops.encode_g1(o, d, t, 6.0)
        prec = ops.PRECISIONS[args.precision]
        nbytes = L.nerf_g1_mlp_workspace_bytes(b * s, 3, 3, prec, 1)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        sigma = torch.empty(b, s, device=dev); rgb = torch.empty(b, s, 3, device=dev)
        dsig = torch.randn(b, s, device=dev); drgb = torch.randn(b, s, 3, device=dev)
        grads = torch.zeros_like(flat)

        def mlp_only(i):
            _lib.check(L.nerf_g1_mlp_fwd(P(flat), P(ex), P(ed), b * s, s, 3, 3, prec, 1, P(sigma), P(rgb), P(ws), nbytes, _lib.stream()), "fwd")
            _lib.check(L.nerf_g1_mlp_bwd(P(flat), P(ex), P(ed), b * s, s, 3, 3, prec, P(dsig), P(drgb), P(grads), P(ws), nbytes, _lib.stream()), "bwd")
        for i in range(3):
            mlp_only(i)
        ms_mlp = float(np.mean(timed_steps(mlp_only, max(K // 2, 3), flush)))
        mlp_kernel_name = "G1 MLP fwd+bwd (nerf_g1_mlp_fwd + nerf_g1_mlp_bwd, FFMA tiles)"
    mlp_tflops = FLOP_PER_POINT_FWD_BWD * b * s / (ms_mlp * 1e-3) / 1e12

    # ---- compositing fwd+bwd alone at a sweep size larger than L2 (HBM roofline) ----
    cb = 1 << 20
    csig = torch.randn(cb, s, device=dev).relu_() * 10; crgb = torch.randn(cb, s, 3, device=dev)
    ct = (torch.rand(cb, s, device=dev) * 4 + 2).sort(dim=-1)[0]
    co = torch.zeros(cb, 3, device=dev); cd = torch.randn(cb, 3, device=dev)
    cpix = torch.empty(cb, 3, device=dev); gp = torch.randn(cb, 3, device=dev)
    dsg = torch.empty_like(csig); drg = torch.empty_like(crgb)

    def comp_fwd(i):
        _lib.check(_lib.lib().nerf_composite_fwd(P(csig), P(crgb), None, P(co), P(cd), P(ct), cb, s, P(cpix), None, None, None, _lib.stream()), "cf")

    def comp_bwd(i):
        _lib.check(_lib.lib().nerf_composite_bwd(P(csig), P(crgb), None, P(co), P(cd), P(ct), cb, s, P(gp), None, None, None, P(cpix), None, None, P(dsg), P(drg), _lib.stream()), "cb")
    comp_fwd(0); comp_bwd(0)
    ms_cf = float(np.mean(timed_steps(comp_fwd, 5, flush)))
    ms_cb = float(np.mean(timed_steps(comp_bwd, 5, flush)))
    comp_fwd_gbs = COMPOSITE_FWD_BYTES_PER_RAY * cb / (ms_cf * 1e-3) / 1e9
    comp_bwd_gbs = COMPOSITE_BWD_BYTES_PER_RAY * cb / (ms_cb * 1e-3) / 1e9
    # the stratified sampler alone at the same size: 4 S bytes per ray written (+ 4 S read with host-supplied uniforms)
    cu = torch.rand(cb, s, device=dev)

    def samp_u(i):
        _lib.check(_lib.lib().nerf_sample_linear(NEAR, FAR, None, s, cb, P(cu), 1, 6, 0, P(ct), _lib.stream()), "su")

    def samp_philox(i):
        _lib.check(_lib.lib().nerf_sample_linear(NEAR, FAR, None, s, cb, None, 1, 6, 0, P(ct), _lib.stream()), "sp")
    samp_u(0); samp_philox(0)
    ms_su = float(np.mean(timed_steps(samp_u, 5, flush)))
    ms_sp = float(np.mean(timed_steps(samp_philox, 5, flush)))
    samp_u_gbs = 8 * s * cb / (ms_su * 1e-3) / 1e9
    samp_p_gbs = 4 * s * cb / (ms_sp * 1e-3) / 1e9
    del csig, crgb, ct, dsg, drg, cu

    # ---- render throughput: 800x800 x 64 (fused bf16 path) or 200x200 (fp32 path) through render_image ----
    from oracle.g1 import pose_spherical, TINY_NERF_FOCAL
    pose = pose_spherical(30.0, -30.0, 4.0311).to(dev)
    rh = rw = 800 if args.precision == "bf16" else 200
    rfocal = TINY_NERF_FOCAL * rh / 100
    with torch.no_grad():
        model.render_image(pose[None, :3, 3], pose[None, :3, :3], rh, rw, rfocal, NEAR, FAR, SAMPLES)
        ms_r = timed_steps(lambda i: model.render_image(pose[None, :3, 3], pose[None, :3, :3], rh, rw,
                                                        rfocal, NEAR, FAR, SAMPLES), 3, flush)
    render_rps = rh * rw / (float(np.mean(ms_r)) * 1e-3)
    render_tflops = render_rps * SAMPLES * FLOP_PER_POINT_FWD / 1e12
    # D3: 64 coarse + 128 fine samples (hierarchical inverse-CDF pass, 192-point fine evaluation)
    d3 = None
    if args.precision == "bf16":
        with torch.no_grad():
            model.render_image(pose[None, :3, 3], pose[None, :3, :3], rh, rw, rfocal, NEAR, FAR, SAMPLES, num_fine=128)
            ms_d3 = timed_steps(lambda i: model.render_image(pose[None, :3, 3], pose[None, :3, :3], rh, rw,
                                                             rfocal, NEAR, FAR, SAMPLES, num_fine=128), 2, flush)
        d3_rps = rh * rw / (float(np.mean(ms_d3)) * 1e-3)
        d3 = {"value": d3_rps, "unit": "rays/s", "workload": f"{rh}x{rw}, 64 coarse + 128 fine samples (256 MLP evaluations per ray)",
              "mlp_tflops": d3_rps * 256 * FLOP_PER_POINT_FWD / 1e12, "frac_tensor": d3_rps * 256 * FLOP_PER_POINT_FWD / 1e12 / peaks["tensor_sustained"]}

    # ---- HEAD generation (fp32 FFMA path): render_image 200x200 x 64 fenceposts (63 intervals per ray) ----
    from nerf_b200.head import HeadNeRF
    torch.manual_seed(3)
    hm = HeadNeRF().to(dev)
    hh = 200
    with torch.no_grad():
        hm.render_image(pose[None, :3, 3], pose[None, :3, :3], hh, hh, 224.0, SAMPLES)
        ms_h = timed_steps(lambda i: hm.render_image(pose[None, :3, 3], pose[None, :3, :3], hh, hh, 224.0, SAMPLES), 3, flush)
    head_rps = hh * hh / (float(np.mean(ms_h)) * 1e-3)
    head_tflops = head_rps * (SAMPLES - 1) * 601_088 / 1e12
    # ... its fp32 training step through the public API (render_rays -> loss.backward(): composite + trunk backward
    # kernels of head.cu; torch.optim.Adam on the 22 tensors), 4096 rays x 64 fenceposts
    hopt = torch.optim.Adam(hm.parameters(), lr=1e-4)
    hg = torch.Generator(device="cpu").manual_seed(8)
    ho = (torch.randn(RAYS, 3, generator=hg) * 2).to(dev); hd = torch.randn(RAYS, 3, generator=hg).to(dev)
    htgt = torch.rand(RAYS, 3, generator=hg).to(dev); hlab = torch.randint(0, 50, (RAYS,), generator=hg).to(dev)

    def head_step(i):
        hopt.zero_grad(set_to_none=True)
        himg, hseg = hm.render_rays(ho, hd, SAMPLES, randomly_sample=True, density_noise_std=1.0, seed=i)
        hl = ((himg[:, 0] - htgt) ** 2).mean() + torch.nn.functional.nll_loss(hseg[:, 0], hlab)
        hl.backward(); hopt.step()
    head_step(0)
    ms_ht = timed_steps(head_step, 3, flush)
    head_train_rps = RAYS / (float(np.mean(ms_ht)) * 1e-3)
    # ... and its tcgen05 bf16 path (head_tc.cu): 800x800
    torch.manual_seed(3)
    hb = HeadNeRF(precision="bf16").to(dev)
    hbh = 800
    with torch.no_grad():
        hb.render_image(pose[None, :3, 3], pose[None, :3, :3], hbh, hbh, 896.0, SAMPLES)
        ms_hb = timed_steps(lambda i: hb.render_image(pose[None, :3, 3], pose[None, :3, :3], hbh, hbh, 896.0, SAMPLES), 3, flush)
    headb_rps = hbh * hbh / (float(np.mean(ms_hb)) * 1e-3)
    headb_tflops = headb_rps * (SAMPLES - 1) * 601_088 / 1e12

    # ---- C4 / D4: the same step with latent-coded inputs (8 scenes, states_x in [-1,1]^4 => density_inputs 7) ----
    c4_ms = None
    c4_render = None
    if args.precision == "bf16":
        torch.manual_seed(5)
        cm = NeRF(normalize_position=6.0, density_inputs=7, precision="bf16").to(dev)
        sx1 = torch.tensor([[0.3, -0.2, 0.7, -0.9]], device=dev)         # validation render of train_conditional_nerf.py:141-153
        with torch.no_grad():
            cm.render_image(pose[None, :3, 3], pose[None, :3, :3], rh, rw, rfocal, NEAR, FAR, SAMPLES, states_x=sx1)
            ms_c4r = timed_steps(lambda i: cm.render_image(pose[None, :3, 3], pose[None, :3, :3], rh, rw, rfocal, NEAR, FAR, SAMPLES,
                                                           states_x=sx1), 3, flush)
        c4_rps = rh * rw / (float(np.mean(ms_c4r)) * 1e-3)
        c4_render = {"value": c4_rps, "unit": "rays/s", "workload": f"conditional {rh}x{rw} x {SAMPLES} samples render_image(states_x=[1,4])",
                     "mlp_tflops": c4_rps * SAMPLES * 1_343_488 / 1e12}        # G1 + 4 latents: 671 744 MAC per point
        ctr = Trainer(cm, lr=1e-4, near=NEAR, far=FAR, num_samples=SAMPLES, density_noise_std=1.0, randomly_sample=True, seed=4321)
        g4 = torch.Generator().manual_seed(4)
        scenes = torch.rand(8, 4, generator=g4) * 2 - 1
        c4_pool = [scenes[torch.randint(0, 8, (RAYS,), generator=g4)].to(dev) for _ in range(n_pool)]

        def step_c4(i):
            o_, d_, tgt_ = pool_d[i % n_pool]
            return ctr.step(o_, d_, tgt_, states_x=c4_pool[i % n_pool])
        for i in range(3):
            step_c4(i)
        barrier()
        c4_ms = float(np.sum(timed_steps(step_c4, min(K, 50), flush))) / min(K, 50)

    # max over ranks of the timed totals
    if world > 1:
        tt = torch.tensor([total_ms, total_e2e, c4_ms or 0.0], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms, total_e2e, c4_max = tt.tolist()
        c4_ms = c4_max if c4_ms is not None else None
    value = world * RAYS * K / (total_ms * 1e-3)
    e2e = world * RAYS * K / (total_e2e * 1e-3)

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            rps, secs = cpu_train_step_rays_per_s(1024, 8, warmup=1)
            cpu = {"value": rps, "unit": "rays/s", "cores": os.cpu_count(), "kind": "port",
                   "sample": f"oracle G1 train step (fwd+bwd+Adam), 1024 rays x {SAMPLES} samples, median of 8 ({secs:.1f} s)"}
        tensor_peak = peaks["tensor_sustained"]
        line = {
            "metric": "rays/sec (train fwd+bwd)", "value": value, "unit": "rays/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32" if args.precision == "fp32" else "bf16", "data": "synthetic",
            "config": {"workload": "tiny NeRF train step fwd+bwd+Adam, 4096-ray batches x 64 samples (G1, nerf.pth), per GPU",
                       "rays_per_gpu": RAYS, "samples": SAMPLES, "precision": args.precision,
                       "l2": "256 MiB flush between timed steps", "parallelism": f"dp{world}"},
            "e2e": {"value": e2e, "unit": "rays/s", "h2d_bytes_per_step": RAYS * 9 * 4, "d2h_bytes_per_step": 4},
            "gpu_launches": trainer.kernel_launches * K,
            "clocks": clk.summary(),
            "roofline": dominant_roofline(kernels, peaks) if kernels else
                        {"bound": "tensor", "kernel": mlp_kernel_name, "achieved": mlp_tflops, "peak": tensor_peak, "unit": "TFLOP/s",
                         "frac": mlp_tflops / tensor_peak, "traffic": None, "peak_source": peaks["src"] + " bf16 sustained", "ms": ms_mlp},
            "roofline_mlp_step": {"bound": "tensor", "kernel": mlp_kernel_name, "achieved": mlp_tflops, "peak": tensor_peak, "unit": "TFLOP/s",
                                  "frac": mlp_tflops / tensor_peak, "ms": ms_mlp},
            "roofline_hbm": {"composite_fwd": {"achieved": comp_fwd_gbs, "frac": comp_fwd_gbs / peaks["hbm"], "ms": ms_cf},
                             "composite_bwd": {"achieved": comp_bwd_gbs, "frac": comp_bwd_gbs / peaks["hbm"], "ms": ms_cb},
                             "sample_host_u": {"achieved": samp_u_gbs, "frac": samp_u_gbs / peaks["hbm"], "ms": ms_su},
                             "sample_philox": {"achieved": samp_p_gbs, "frac": samp_p_gbs / peaks["hbm"], "ms": ms_sp,
                                               "note": "fma-pipe bound: 20 IMAD.WIDE per 4 samples (Philox4x32-10)"},
                             "peak": peaks["hbm"], "unit": "GB/s", "rays": cb, "samples": s},
            "render": {"value": render_rps, "unit": "rays/s", "workload": f"{rh}x{rw} x {SAMPLES} samples render_image",
                       "mlp_tflops": render_tflops, "frac_tensor": render_tflops / tensor_peak},
            "train_c4": None if c4_ms is None else {
                "value": world * RAYS / (c4_ms * 1e-3), "unit": "rays/s", "ms_per_step": c4_ms,
                "workload": "conditional train step (8 scenes, 4 latents per ray: block_0.0.weight [256,140]), 4096 rays x 64 samples per GPU"},
            "render_d3": d3,
            "render_c4": c4_render,
            "kernels": kernels,
            "head_render": {"value": head_rps, "unit": "rays/s", "workload": f"HEAD (mip-NeRF) {hh}x{hh} x {SAMPLES} fenceposts render_image, fp32 FFMA",
                            "mlp_tflops": head_tflops},
            "head_train_fp32": {"value": head_train_rps, "unit": "rays/s", "ms_per_step": float(np.mean(ms_ht)),
                                "workload": f"HEAD (mip-NeRF) train step fwd+bwd+Adam, {RAYS} rays x {SAMPLES} fenceposts, fp32 FFMA (head.cu), colour MSE + segmentation NLL",
                                "mlp_tflops": head_train_rps * (SAMPLES - 1) * 1_754_112 / 1e12},
            "head_render_bf16": {"value": headb_rps, "unit": "rays/s", "workload": f"HEAD (mip-NeRF) {hbh}x{hbh} x {SAMPLES} fenceposts render_image, tcgen05 bf16 (head_tc_kernel)",
                                 "mlp_tflops": headb_tflops, "frac_tensor": headb_tflops / tensor_peak},
            "loss": loss_val,
        }
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="bf16", choices=["fp32", "bf16"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
