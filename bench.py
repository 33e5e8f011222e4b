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


NCU_TRAIN = ("profiles/r02_train_tc_ncu.json", "profiles/r01_train_tc_v10_ncu.json")     # newest first
NCU_RENDER = ("profiles/r02_render_tc_ncu.json", "profiles/r01_render_fwd_v5_ncu.json")


def ncu_counters(paths, names):
    """Per-launch ncu counters of the committed `ncu --set full` summaries under profiles/ (scripts/ncu_summary.py, taken at
    this workload size): {kernel: {dram_bytes, tensor_pipe_pct, source}}; {} when absent."""
    out = {}
    for rel in paths:
        path = os.path.join(ROOT, rel)
        if not os.path.exists(path):
            continue
        for k in json.load(open(path))["kernels"]:
            for name in names:
                if name in k["kernel"] and name not in out and "dram_bytes" in k:
                    out[name] = {"dram_bytes": k["dram_bytes"], "tensor_pipe_pct": k.get("tensor_pipe_pct"), "source": rel}
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


def bench_config(world):
    """The `config` object, identical in both arms (the reference arm runs the same 4096 x 64 batch per step)."""
    return {"workload": "tiny NeRF train step fwd+bwd+Adam, 4096-ray batches x 64 samples (G1, nerf.pth), per GPU",
            "rays_per_gpu": RAYS, "samples": SAMPLES,
            "l2": "no flush needed: every step streams its own 2.7 GB activation workspace (>> 126 MB L2) and K steps are "
                  "timed back to back in ONE CUDA-event bracket (launch gaps included); the CPU arm has no L2 policy",
            "parallelism": f"dp{world}"}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path (the oracle port: the G1 source is not
    in the reference tree) with all host threads, on the SAME workload: 4096 rays x 64 samples per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.perf_counter()
    rps, _ = cpu_train_step_rays_per_s(RAYS, args.steps, args.warmup)
    ms = 1e3 * RAYS / rps
    line = {"impl": "reference", "metric": "rays/sec (train fwd+bwd)", "value": rps, "unit": "rays/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": bench_config(args.gpus),
            "cpu_baseline": {"value": rps, "unit": "rays/s", "cores": os.cpu_count(), "kind": "port",
                             "sample": f"the full workload: {RAYS} rays x {SAMPLES} samples per step, median of {args.steps} steps"},
            "e2e": {"value": rps, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": time.perf_counter() - t0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ our arm -----
def timed_steps(fn, steps, flush=None):
    """Per-call CUDA-event brackets on the current stream (kernel-level timings: each call alone, L2 flushed by a
    256 MiB write outside the bracket when `flush` is given)."""
    evs = []
    for i in range(steps):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(i); b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    return [a.elapsed_time(b) for a, b in evs]


def timed_region(fn, steps):
    """EXACTLY `steps` calls back to back inside ONE CUDA-event bracket (synchronised on both sides by the caller's
    barrier): device time of the whole region, launch gaps included.  Returns (device ms, wall ms)."""
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.perf_counter()
    a.record()
    for i in range(steps):
        fn(i)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b), 1e3 * (time.perf_counter() - w0)


def safe(section, fn):
    """Secondary measurements never take the headline line down with them."""
    try:
        return fn()
    except Exception as e:      # noqa: BLE001
        return {"error": f"{section}: {type(e).__name__}: {e}"[:300]}


def run_ours(args):
    import torch.distributed as dist
    from nerf_b200 import _lib, ops, parallel
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
    sd = nerf_pth_state_dict()
    K, W = args.steps, args.warmup
    b, s = RAYS, SAMPLES

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(*vals):
        if world == 1:
            return list(vals)
        t = torch.tensor(vals, device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()

    clk = ClockSampler(local).__enter__()                 # before the warm-up: the timed region at --steps 20 is ~30 ms

    model = NeRF(normalize_position=6.0, precision=args.precision).to(dev)
    model.load_state_dict(sd)
    trainer = Trainer(model, lr=1e-4, near=NEAR, far=FAR, num_samples=SAMPLES, density_noise_std=1.0,
                      randomly_sample=True, seed=1234)
    # a pool of distinct batches, resident in HBM (value) and in pinned host memory (e2e)
    n_pool = 8
    pool_h = []
    for i in range(n_pool):
        o, d = synthetic_batch(RAYS, seed=1000 * rank + i)
        tgt = torch.rand(RAYS, 3, generator=torch.Generator().manual_seed(7 + 100 * rank + i))
        pool_h.append(torch.stack([o, d, tgt]).pin_memory())      # one pinned [3, RAYS, 3] block per batch: one H2D copy
    pool_d = [tuple(x.to(dev)) for x in pool_h]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def step_resident(i):
        o, d, tgt = pool_d[i % n_pool]
        return trainer.step(o, d, tgt)

    # e2e: every step's batch comes from pinned HOST memory and every step's loss goes back to the host, through the repo's
    # input pipeline (train.HostBatchPipeline: the copy of batch i+1 runs on a copy stream under step i)
    from nerf_b200.train import HostBatchPipeline
    pipe = HostBatchPipeline(dev, like=pool_h[0])
    e2e_state = {"n": 0}

    def e2e_begin(n):
        e2e_state["n"] = n
        pipe.submit(pool_h[0])

    def step_e2e(i):
        if i + 1 < e2e_state["n"]:
            pipe.submit(pool_h[(i + 1) % n_pool])
        o, d, tgt = pipe.get()                                     # rays_o, rays_d, target pixels: views of one [3, RAYS, 3] buffer
        loss = trainer.step(o, d, tgt)
        pipe.done(loss, ring=trainer.loss_ring)

    # ------------------------------------------------------------------ kernels timed alone through the C ABI ----
    # Taken TWICE: here, after three steps of warm-up and before any long region (boost clocks: the state the BURST peak of
    # MEASURED_PEAKS.json is measured in), and again after the headline region (power-capped: the SUSTAINED state).  The same
    # kernels differ by 8-12 % between the two states (0.382 / 0.513 / 0.438 ms against 0.436 / 0.575 / 0.448 ms), with or
    # without an L2 flush in front of them: their per-launch traffic is 1.4-2.9 GB against 126 MB of L2.
    o, d, tgt = pool_d[0]
    flat = model.flat_parameters()
    P = _lib.ptr
    L = _lib.lib()
    kernel_ms_burst = None
    if args.precision == "bf16":
        for i in range(3):
            step_resident(i)
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

        def time_kernels(n_it):
            return (float(np.mean(timed_steps(k_fwd, n_it, flush))), float(np.mean(timed_steps(k_bwd(1), n_it, flush))),
                    float(np.mean(timed_steps(k_bwd(3), n_it, flush))))
        mlp_only(0)
        torch.cuda.synchronize()
        kernel_ms_burst = time_kernels(5)

    # ------------------------------------------------------------------ headline: K steps, one bracket ----
    for i in range(W):
        step_resident(i)
    barrier()
    total_ms, wall_ms = timed_region(step_resident, K)
    barrier()
    n_w = max(W // 2, 1)
    e2e_begin(n_w)
    for i in range(n_w):
        step_e2e(i)
    barrier()

    def e2e_region(i):                                             # the first batch's copy is issued INSIDE the timed region
        if i == 0:
            e2e_begin(K)
        step_e2e(i)
        if i == K - 1:                                             # the closing event waits for the last loss read-back too
            torch.cuda.current_stream().wait_event(pipe.loss_read)
    total_e2e, wall_e2e = timed_region(e2e_region, K)
    pipe.d2h.synchronize()
    barrier()
    host_losses = pipe.losses()
    loss_val = float(host_losses[-1].item())
    assert host_losses.numel() == n_w + K and bool(torch.isfinite(host_losses).all())
    launches_per_step = trainer.kernel_launches

    # DP correctness on the hardware: parameters identical on every rank after the timed steps, and the
    # all-reduced gradient of one more step equal to the 1-GPU gradient of the CONCATENATED batch
    dp_check = None
    allreduce_us = None
    if world > 1:
        def _dp():
            flat = model.flat_parameters()
            lo_, hi_ = flat.clone(), flat.clone()
            dist.all_reduce(lo_, op=dist.ReduceOp.MIN); dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
            param_spread = float((hi_ - lo_).abs().max().item())
            # every rank rebuilds ALL ranks' batches (synthetic, seeded by rank) and runs them as ONE batch without the exchange
            o_all = torch.cat([synthetic_batch(512, seed=5000 + r)[0] for r in range(world)]).to(dev)
            d_all = torch.cat([synthetic_batch(512, seed=5000 + r)[1] for r in range(world)]).to(dev)
            t_all = torch.cat([torch.rand(512, 3, generator=torch.Generator().manual_seed(900 + r)) for r in range(world)]).to(dev)
            u_all = torch.cat([torch.rand(512, s, generator=torch.Generator().manual_seed(950 + r)) for r in range(world)]).to(dev)
            sl = slice(512 * rank, 512 * (rank + 1))
            tr_dp = Trainer(model, num_samples=s, density_noise_std=0.0, precision=args.precision, fused_exchange=False)
            tr_dp.step(o_all[sl].contiguous(), d_all[sl].contiguous(), t_all[sl].contiguous(), u=u_all[sl].contiguous(), apply_update=False)
            g_dp = tr_dp.flat_grad.clone() / world                          # what Adam sees (grad_scale = 1/world)
            tr_one = Trainer(model, num_samples=s, density_noise_std=0.0, precision=args.precision, fused_exchange=False)
            tr_one.world = 1                                               # same kernels, no exchange
            tr_one.step(o_all, d_all, t_all, u=u_all, apply_update=False)
            rel = float(((g_dp - tr_one.flat_grad).norm() / tr_one.flat_grad.norm()).item())
            return {"param_max_abs_spread_across_ranks": param_spread, "grad_rel_l2_vs_single_gpu_concatenated_batch": rel,
                    "rays_per_rank": 512, "ok": bool(param_spread == 0.0 and rel <= 1e-4),
                    "tolerance": "parameters: 0; gradient: 1e-4 rel-L2 (fp32 summation order of N partial sums; measured 4.3e-6 at N = 2, 2.3e-5 at N = 8)"}
        dp_check = safe("dp_check", _dp)

        def _ar():
            g = torch.zeros_like(trainer.flat_grad)
            for _ in range(5):
                dist.all_reduce(g)
            barrier()
            ms = timed_steps(lambda i: dist.all_reduce(g), 20)
            return {"us": 1e3 * float(np.median(ms)), "bytes": g.numel() * 4, "what": "ncclAllReduce(sum, fp32) of the flat gradient, alone"}
        allreduce_us = safe("allreduce", _ar)

    # ------------------------------------------------------------------ the same kernels again, after the long region ----
    kernels = {}
    roofline = None
    if args.precision == "bf16":
        for i in range(3):
            mlp_only(i)
        n_it = min(max(K // 2, 5), 20)
        ms_fwd_s, ms_dgrad_s, ms_wg_s = time_kernels(n_it)
        ms_mlp, _ = timed_region(mlp_only, n_it); ms_mlp /= n_it
        ms_fwd, ms_dgrad, ms_wg = kernel_ms_burst
        n_tiles = (b * s + 127) // 128
        wgrad_bytes = n_tiles * ((10 + 9 + 2) * 65536 + 3 * 16384 + 2 * 4096)     # dz + unit inputs + head inputs + encodings + dh
        act_bytes = n_tiles * 65536                                                # one unit's [128 x 256] bf16 slabs, all tiles
        fwd_bytes = 10 * act_bytes + 2 * n_tiles * 16384 + 10 * n_tiles * 128 * 8 + b * s * 20   # activations, encodings, stats, t/sigma/rgb
        dgrad_bytes = 20 * act_bytes + 10 * n_tiles * 128 * 8 + n_tiles * 4096 + b * s * 16      # n_l in, dz_l out, stats, dh image, d_sigma/d_rgb
        ncu = ncu_counters(NCU_TRAIN, ("g1_fused_tc_kernel", "g1_bwd_tc_kernel", "g1_wgrad_tc_kernel"))
        io_bytes_8d = b * 24 + b * s * 20                  # what the MLP must move -- rays in, (t, sigma, rgb) out

        def kern(name, ms, ms_s, flop, design_bytes):
            tf, gbs = flop / ms / 1e9, design_bytes / ms / 1e6
            c = ncu.get(name, {})
            return {"ms": ms, "flop": flop,
                    "after_long_run": {"ms": ms_s, "tflops": flop / ms_s / 1e9, "frac_of_sustained_peak": flop / ms_s / 1e9 / peaks["tensor_sustained"],
                                       "note": "the same launch timed alone after the headline region (power-capped clocks)"},
                    # the MLP kernels are scored on the TENSOR roofline; a kernel timed alone gets the BURST peak
                    "tensor": {"achieved": tf, "unit": "TFLOP/s", "peak": peaks["tensor_burst"], "frac": tf / peaks["tensor_burst"],
                               "frac_of_sustained": tf / peaks["tensor_sustained"], "tensor_pipe_pct_ncu": c.get("tensor_pipe_pct")},
                    # the HBM view of THIS design: the bf16 activation stash is traffic the design chose, not algorithmic bytes
                    "hbm_design": {"achieved": gbs, "unit": "GB/s", "peak": peaks["hbm"], "frac": gbs / peaks["hbm"], "stash_bytes": design_bytes,
                                   "traffic_ncu": c.get("dram_bytes"), "ncu_source": c.get("source")}}
        kernels = {
            "g1_fused_tc_kernel<2,train>": kern("g1_fused_tc_kernel", ms_fwd, ms_fwd_s, FLOP_PER_POINT_FWD * b * s, fwd_bytes),
            "g1_bwd_tc_kernel<2>": kern("g1_bwd_tc_kernel", ms_dgrad, ms_dgrad_s, 2 * 9 * 65536 * b * s, dgrad_bytes),
            "g1_wgrad_tc_kernel": kern("g1_wgrad_tc_kernel", ms_wg, ms_wg_s, 2 * (9 * 65536 + 3 * 16384 + 10 * 256 + 2 * 1024) * b * s, wgrad_bytes),
        }
        name, k = max(kernels.items(), key=lambda kv: kv[1]["ms"])
        traffic = k["hbm_design"]["traffic_ncu"]
        roofline = {"bound": "tensor", "kernel": name, "achieved": k["tensor"]["achieved"], "peak": peaks["tensor_burst"], "unit": "TFLOP/s",
                    "frac": k["tensor"]["frac"], "traffic": traffic, "ms": k["ms"],
                    "peak_source": peaks["src"] + " bf16 burst (kernel timed alone, before the long region); SURVEY 8(d): MLP kernels on the tensor roofline",
                    "after_long_run": k["after_long_run"],
                    "tensor_pipe_pct_ncu": k["tensor"]["tensor_pipe_pct_ncu"], "flop_per_launch": k["flop"],
                    "algorithmic_bytes_8d": io_bytes_8d,
                    "wasted_traffic_x": None if not traffic else traffic / io_bytes_8d}
        total_traffic = sum((v["hbm_design"]["traffic_ncu"] or v["hbm_design"]["stash_bytes"]) for v in kernels.values())
        roofline_hbm_design = {"note": "HBM view of the activation-stash design: bytes the three kernels move per step against the copy peak",
                               "bytes_per_step": total_traffic, "floor_ms_at_copy_peak": total_traffic / peaks["hbm"] / 1e6,
                               "vs_8d_bytes_x": total_traffic / io_bytes_8d,
                               "kernels": {kn: v["hbm_design"] for kn, v in kernels.items()}}
        mlp_kernel_name = "G1 MLP fwd+bwd on tcgen05 (" + " + ".join(kernels) + " + unfold)"
    else:
        t = ops.sample_linear(NEAR, FAR, s, b, dev, jitter=True, seed=1)
        ex, ed, _ = ops.encode_g1(o, d, t, 6.0)
        nbytes = L.nerf_g1_mlp_workspace_bytes(b * s, 3, 3, 1)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        sigma = torch.empty(b, s, device=dev); rgb = torch.empty(b, s, 3, device=dev)
        dsig = torch.randn(b, s, device=dev); drgb = torch.randn(b, s, 3, device=dev)
        grads = torch.zeros_like(flat)

        def mlp_only(i):
            _lib.check(L.nerf_g1_mlp_fwd(P(flat), P(ex), P(ed), b * s, s, 3, 3, 1, P(sigma), P(rgb), P(ws), nbytes, _lib.stream()), "fwd")
            _lib.check(L.nerf_g1_mlp_bwd(P(flat), P(ex), P(ed), b * s, s, 3, 3, P(dsig), P(drgb), P(grads), P(ws), nbytes, _lib.stream()), "bwd")
        for i in range(3):
            mlp_only(i)
        ms_mlp = float(np.mean(timed_steps(mlp_only, max(K // 2, 3), flush)))
        mlp_kernel_name = "G1 MLP fwd+bwd (nerf_g1_mlp_fwd + nerf_g1_mlp_bwd, FFMA tiles)"
        roofline_hbm_design = None
    mlp_tflops = FLOP_PER_POINT_FWD_BWD * b * s / (ms_mlp * 1e-3) / 1e12
    if roofline is None:
        roofline = {"bound": "tensor", "kernel": mlp_kernel_name, "achieved": mlp_tflops, "peak": peaks["tensor_sustained"], "unit": "TFLOP/s",
                    "frac": mlp_tflops / peaks["tensor_sustained"], "traffic": None, "peak_source": peaks["src"] + " bf16 sustained", "ms": ms_mlp}

    # ------------------------------------------------------------------ the reference loop through the public API ----
    def _api_loop():
        m2 = NeRF(normalize_position=6.0, precision=args.precision).to(dev); m2.load_state_dict(sd)
        opt = torch.optim.Adam(m2.parameters(), lr=1e-4)

        def it(i):
            o, d, tgt = pool_d[i % n_pool]
            px = m2.render_rays(o, d, NEAR, FAR, SAMPLES, randomly_sample=True, density_noise_std=1.0)
            opt.zero_grad()
            ((px - tgt) ** 2).mean().backward()
            opt.step()
        for i in range(3):
            it(i)
        torch.cuda.synchronize()
        n = min(K, 50)
        ms, wall = timed_region(it, n)
        torch.cuda.synchronize()
        h0 = time.perf_counter()                               # host time to ISSUE a step into an empty queue (no back-pressure)
        for i in range(10):
            it(i)
        host_ms = 1e2 * (time.perf_counter() - h0)
        torch.cuda.synchronize()
        return {"value": RAYS * n / (ms * 1e-3), "unit": "rays/s", "ms_per_step": ms / n, "wall_ms_per_step": wall / n, "steps": n,
                "host_issue_ms_per_step": host_ms,
                "note": "host-bound when host_issue_ms_per_step ~ ms_per_step: the loop's Python side (autograd over 44 parameter "
                        "tensors, torch.optim.Adam's foreach kernels) takes as long as the device's 1.46 ms of kernels",
                "what": "example.ipynb cell 8 / train_conditional_nerf.py:120-135 verbatim: pixels = nerf.render_rays(...); "
                        "loss.backward(); torch.optim.Adam.step() -- per GPU, no gradient exchange",
                "tcgen05_calls": {k: v for k, v in ops.COUNTERS.items() if "tc" in k}}
    api_loop = safe("api_loop", _api_loop)

    # ------------------------------------------------------------------ the other precision of config 2 ----
    def _other_precision():
        other = "fp32" if args.precision == "bf16" else "bf16"
        m3 = NeRF(normalize_position=6.0, precision=other).to(dev); m3.load_state_dict(sd)
        tr3 = Trainer(m3, lr=1e-4, near=NEAR, far=FAR, num_samples=SAMPLES, density_noise_std=1.0, randomly_sample=True, seed=1234,
                      fused_exchange=False)
        tr3.world = 1

        def it(i):
            o, d, tgt = pool_d[i % n_pool]
            tr3.step(o, d, tgt)
        for i in range(2):
            it(i)
        torch.cuda.synchronize()
        n = min(K, 10) if other == "fp32" else min(K, 50)
        ms, _ = timed_region(it, n)
        return {"value": RAYS * n / (ms * 1e-3), "unit": "rays/s", "ms_per_step": ms / n, "steps": n, "precision": other, "per_gpu": True,
                "contraction": "fp32 FFMA tiles (mlp_fp32_*.cuh)" if other == "fp32" else "tcgen05 kind::f16",
                "mlp_tflops": RAYS * SAMPLES * FLOP_PER_POINT_FWD_BWD * n / (ms * 1e-3) / 1e12}
    other_prec = safe("other_precision", _other_precision)

    # ------------------------------------------------------------------ fp32-grade step on the tensor cores (bf16x3) ----
    def _x3_train():
        m4 = NeRF(normalize_position=6.0, precision="bf16x3").to(dev); m4.load_state_dict(sd)
        tr4 = Trainer(m4, lr=1e-4, near=NEAR, far=FAR, num_samples=SAMPLES, density_noise_std=1.0, randomly_sample=True, seed=1234,
                      fused_exchange=False)
        tr4.world = 1

        def it(i):
            o, d, tgt = pool_d[i % n_pool]
            tr4.step(o, d, tgt)
        for i in range(3):
            it(i)
        torch.cuda.synchronize()
        n = min(K, 20)
        ms, _ = timed_region(it, n)
        z4 = tr4._tc_buffers(RAYS, SAMPLES, dev)
        flat4 = m4.flat_parameters(); P4 = _lib.ptr; L4 = _lib.lib(); g4 = torch.zeros_like(flat4)
        o, d, tgt = pool_d[0]
        lad = m4._ladder(NEAR, FAR, SAMPLES, dev)

        def k_fwd(i):
            _lib.check(L4.nerf_g1_train_fwd_x3(P4(z4["packed"]), P4(o), P4(d), RAYS, SAMPLES, NEAR, FAR, P4(lad), None, None, 1, 1, 0, 6.0, 1,
                                               P4(z4["t"]), P4(z4["sigma"]), P4(z4["rgb"]), P4(z4["ws"]), z4["n"], _lib.stream()), "fwd_x3")

        def k_bwd(stage):
            def f(i):
                _lib.check(L4.nerf_g1_train_bwd_x3(P4(z4["packed"]), P4(z4["packed_bwd"]), P4(flat4), P4(z4["d_sigma"]), P4(z4["d_rgb"]), RAYS, SAMPLES,
                                                   P4(g4), P4(z4["ws"]), z4["n"], stage, _lib.stream()), "bwd_x3")
            return f
        ms_f = float(np.mean(timed_steps(k_fwd, 5, flush)))
        ms_d = float(np.mean(timed_steps(k_bwd(1), 5, flush)))
        ms_dw = float(np.mean(timed_steps(k_bwd(2), 5, flush)))
        return {"value": RAYS * n / (ms * 1e-3), "unit": "rays/s", "ms_per_step": ms / n, "steps": n, "precision": "bf16x3", "per_gpu": True,
                "contraction": "tcgen05 kind::f16 with hi/lo operands (3 products per term) in the forward, the data-gradient chain and the "
                               "weight-gradient GEMM (one pass over interleaved hi/lo images); LayerNorm/ReLU forward and backward, heads, "
                               "unfold, Adam in fp32",
                "mlp_tflops": RAYS * SAMPLES * FLOP_PER_POINT_FWD_BWD * n / (ms * 1e-3) / 1e12,
                "kernels_ms": {"g1_fused_x3_kernel<train>": ms_f, "g1_bwd_x3_kernel": ms_d, "g1_wgrad_x3_kernel": ms_dw - ms_d},
                "workspace_bytes": int(z4["n"]),
                "parity": "gradient vs the fp32 path: 7e-4 .. 2e-3 whole-gradient rel-L2 = the exact gradient's own sensitivity to a 1e-5 "
                          "relative change of the forward (profiles/r02_grad_conditioning.txt); bf16: 5e-2 .. 6e-2"}
    x3_train = safe("x3_train", _x3_train) if (world == 1 or args.all_sections) else None

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
    cam_o, cam_r = pose[None, :3, 3].contiguous(), pose[None, :3, :3].contiguous()
    rh = rw = 800 if args.precision == "bf16" else 200
    rfocal = TINY_NERF_FOCAL * rh / 100
    with torch.no_grad():
        model.render_image(cam_o, cam_r, rh, rw, rfocal, NEAR, FAR, SAMPLES)
        ms_r = timed_steps(lambda i: model.render_image(cam_o, cam_r, rh, rw, rfocal, NEAR, FAR, SAMPLES), 3, flush)
    render_rps = rh * rw / (float(np.mean(ms_r)) * 1e-3)
    render_tflops = render_rps * SAMPLES * FLOP_PER_POINT_FWD / 1e12
    rncu = ncu_counters(NCU_RENDER, ("g1_fused_tc_kernel",)).get("g1_fused_tc_kernel", {})

    # ---- fp32-grade render on the tensor cores (precision="bf16x3", hi/lo operand split) beside the fp32 FFMA render ----
    def _render_fp32_grade():
        mx = NeRF(normalize_position=6.0, precision="bf16x3").to(dev); mx.load_state_dict(sd)
        mf = NeRF(normalize_position=6.0, precision="fp32").to(dev); mf.load_state_dict(sd)
        out = {}
        with torch.no_grad():
            for tag, m_, hw_ in (("bf16x3", mx, 800), ("fp32_ffma", mf, 200)):
                f_ = TINY_NERF_FOCAL * hw_ / 100
                m_.render_image(cam_o, cam_r, hw_, hw_, f_, NEAR, FAR, SAMPLES)
                ms_ = float(np.mean(timed_steps(lambda i: m_.render_image(cam_o, cam_r, hw_, hw_, f_, NEAR, FAR, SAMPLES), 2, flush)))
                rps_ = hw_ * hw_ / (ms_ * 1e-3)
                out[tag] = {"value": rps_, "unit": "rays/s", "workload": f"{hw_}x{hw_} x {SAMPLES} samples render_image, one GPU",
                            "mlp_tflops": rps_ * SAMPLES * FLOP_PER_POINT_FWD / 1e12}
            a_ = mx.render_image(cam_o, cam_r, 100, 100, TINY_NERF_FOCAL, NEAR, FAR, SAMPLES)
            b_ = mf.render_image(cam_o, cam_r, 100, 100, TINY_NERF_FOCAL, NEAR, FAR, SAMPLES)
        out["max_abs_pixel_diff_100x100"] = float((a_ - b_).abs().max().item())
        out["speedup"] = out["bf16x3"]["value"] / out["fp32_ffma"]["value"]
        out["note"] = ("fp32-mode accuracy (<= 1e-3 pixels, north star) on tcgen05: every GEMM operand split into two bf16 parts, "
                       "hi*hi + lo*hi + hi*lo with fp32 accumulation (csrc/fused_tc_x3.cu); the training step in this mode: fp32_grade_train")
        return out
    render_fp32_grade = safe("render_fp32_grade", _render_fp32_grade) if (world == 1 or args.all_sections) else None

    # ---- D3 / config 3: 800x800, 64 coarse + 128 fine, RAY-SHARDED over the ranks, frame gathered ----
    d3 = None
    render_sharded = None
    if args.precision == "bf16":
        def _d3():
            kw = dict(num_fine=128)
            total = rh * rw
            with torch.no_grad():
                parallel.render_image_sharded(model, cam_o, cam_r, rh, rw, rfocal, NEAR, FAR, SAMPLES, gather=True, **kw)    # warm-up
                barrier()
                # (i) the shard alone: no collective in the timed region
                lo, hi = parallel.shard_range(total, rank, world)
                ms_sh = float(np.mean(timed_steps(lambda i: model.render_image(cam_o, cam_r, rh, rw, rfocal, NEAR, FAR, SAMPLES,
                                                                                ray_begin=lo, ray_end=hi, **kw), 2)))
                barrier()
                # (ii) shard + all_gather of the [rays/N, 3] pieces into the frame on every rank
                ms_g = timed_steps(lambda i: parallel.render_image_sharded(model, cam_o, cam_r, rh, rw, rfocal, NEAR, FAR, SAMPLES,
                                                                           gather=True, **kw), 2)
                frame = parallel.render_image_sharded(model, cam_o, cam_r, rh, rw, rfocal, NEAR, FAR, SAMPLES, gather=True, **kw)
                ms_g = float(np.mean(ms_g))
                barrier()
                ms_sh, ms_g = max_over_ranks(ms_sh, ms_g)
                equal = None
                if rank == 0:                                        # the gathered frame against the 1-GPU render of the whole frame
                    full = model.render_image(cam_o, cam_r, rh, rw, rfocal, NEAR, FAR, SAMPLES, **kw)
                    equal = {"bit_identical": bool(torch.equal(full, frame)), "max_abs_diff": float((full - frame).abs().max().item())}
                barrier()
            rps = total / (ms_sh * 1e-3)
            return {"value": rps, "unit": "rays/s", "n_gpus": world, "ms_per_frame": ms_sh, "scaling": "strong",
                    "workload": f"{rh}x{rw}, 64 coarse + 128 fine samples (256 MLP evaluations per ray); rank r renders "
                                f"shard_range({total}, r, {world}); no collective in the timed region; time = max over ranks",
                    "with_all_gather": {"value": total / (ms_g * 1e-3), "ms_per_frame": ms_g},
                    "gathered_frame_vs_single_gpu": equal,
                    "mlp_tflops": rps * 256 * FLOP_PER_POINT_FWD / 1e12,
                    "frac_tensor_sustained": rps * 256 * FLOP_PER_POINT_FWD / 1e12 / (world * peaks["tensor_sustained"])}
        render_sharded = safe("render_sharded", _d3)
        if world == 1 and "value" in render_sharded:
            d3 = {"value": render_sharded["value"], "unit": "rays/s", "workload": render_sharded["workload"],
                  "mlp_tflops": render_sharded["mlp_tflops"], "frac_tensor": render_sharded["frac_tensor_sustained"]}

    head = None
    c4 = {"train": None, "render": None}
    if world == 1 or args.all_sections:
        # ---- HEAD generation (fp32 FFMA path): render_image 200x200 x 64 fenceposts (63 intervals per ray) ----
        def _head():
            from nerf_b200.head import HeadNeRF
            torch.manual_seed(3)
            hm = HeadNeRF().to(dev)
            hh = 200
            with torch.no_grad():
                hm.render_image(cam_o, cam_r, hh, hh, 224.0, SAMPLES)
                ms_h = timed_steps(lambda i: hm.render_image(cam_o, cam_r, hh, hh, 224.0, SAMPLES), 3, flush)
            head_rps = hh * hh / (float(np.mean(ms_h)) * 1e-3)
            # ... its fp32 training step through the public API (render_rays -> loss.backward(): composite + trunk backward
            # kernels of head.cu; torch.optim.Adam on the 22 tensors), 4096 rays x 64 fenceposts
            hopt = torch.optim.Adam(hm.parameters(), lr=1e-4)
            hg = torch.Generator(device="cpu").manual_seed(8)
            ho = (torch.randn(RAYS, 3, generator=hg) * 2).to(dev); hd = torch.randn(RAYS, 3, generator=hg).to(dev)
            htgt = torch.rand(RAYS, 3, generator=hg).to(dev); hlab = torch.randint(0, 50, (RAYS,), generator=hg).to(dev)

            def head_step(i):
                hopt.zero_grad(set_to_none=True)
                himg, hseg = hm.render_rays(ho, hd, SAMPLES, randomly_sample=True, density_noise_std=1.0)
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
                hb.render_image(cam_o, cam_r, hbh, hbh, 896.0, SAMPLES)
                ms_hb = timed_steps(lambda i: hb.render_image(cam_o, cam_r, hbh, hbh, 896.0, SAMPLES), 3, flush)
            headb_rps = hbh * hbh / (float(np.mean(ms_hb)) * 1e-3)
            headb_tflops = headb_rps * (SAMPLES - 1) * 601_088 / 1e12
            return {
                "head_render": {"value": head_rps, "unit": "rays/s", "workload": f"HEAD (mip-NeRF) {hh}x{hh} x {SAMPLES} fenceposts render_image, fp32 FFMA",
                                "mlp_tflops": head_rps * (SAMPLES - 1) * 601_088 / 1e12},
                "head_train_fp32": {"value": head_train_rps, "unit": "rays/s", "ms_per_step": float(np.mean(ms_ht)),
                                    "workload": f"HEAD (mip-NeRF) train step fwd+bwd+Adam, {RAYS} rays x {SAMPLES} fenceposts, fp32 FFMA (head.cu), colour MSE + segmentation NLL",
                                    "mlp_tflops": head_train_rps * (SAMPLES - 1) * 1_754_112 / 1e12},
                "head_render_bf16": {"value": headb_rps, "unit": "rays/s", "workload": f"HEAD (mip-NeRF) {hbh}x{hbh} x {SAMPLES} fenceposts render_image, tcgen05 bf16 (head_tc_kernel)",
                                     "mlp_tflops": headb_tflops, "frac_tensor": headb_tflops / peaks["tensor_sustained"]}}
        head = safe("head", _head)

    # ---- C4 / D4: the same step with latent-coded inputs (8 scenes, states_x in [-1,1]^4 => density_inputs 7), data-parallel ----
    if args.precision == "bf16":
        def _c4():
            torch.manual_seed(5)
            cm = NeRF(normalize_position=6.0, density_inputs=7, precision="bf16").to(dev)
            parallel.broadcast_parameters(cm)
            sx1 = torch.tensor([[0.3, -0.2, 0.7, -0.9]], device=dev)         # validation render of train_conditional_nerf.py:141-153
            with torch.no_grad():
                cm.render_image(cam_o, cam_r, rh, rw, rfocal, NEAR, FAR, SAMPLES, states_x=sx1)
                ms_c4r = timed_steps(lambda i: cm.render_image(cam_o, cam_r, rh, rw, rfocal, NEAR, FAR, SAMPLES, states_x=sx1), 3, flush)
            c4_rps = rh * rw / (float(np.mean(ms_c4r)) * 1e-3)
            ctr = Trainer(cm, lr=1e-4, near=NEAR, far=FAR, num_samples=SAMPLES, density_noise_std=1.0, randomly_sample=True, seed=4321)
            g4 = torch.Generator().manual_seed(4 + rank)
            scenes = torch.rand(8, 4, generator=torch.Generator().manual_seed(4)) * 2 - 1
            c4_pool = [scenes[torch.randint(0, 8, (RAYS,), generator=g4)].to(dev) for _ in range(n_pool)]

            def step_c4(i):
                o_, d_, tgt_ = pool_d[i % n_pool]
                return ctr.step(o_, d_, tgt_, states_x=c4_pool[i % n_pool])
            for i in range(3):
                step_c4(i)
            barrier()
            n = min(K, 50)
            ms, _ = timed_region(step_c4, n)
            barrier()
            ms = max_over_ranks(ms)[0]
            return {"train": {"value": world * RAYS * n / (ms * 1e-3), "unit": "rays/s", "ms_per_step": ms / n, "n_gpus": world,
                              "workload": "conditional train step (8 scenes, 4 latents per ray: block_0.0.weight [256,140]), 4096 rays x 64 samples per GPU, data-parallel"},
                    "render": {"value": c4_rps, "unit": "rays/s", "workload": f"conditional {rh}x{rw} x {SAMPLES} samples render_image(states_x=[1,4]), one GPU",
                               "mlp_tflops": c4_rps * SAMPLES * 1_343_488 / 1e12}}        # G1 + 4 latents: 671 744 MAC per point
        r = safe("c4", _c4)
        c4 = r if "train" in r else {"train": r, "render": None}

    # max over ranks of the timed totals
    total_ms, total_e2e, wall_ms, wall_e2e = max_over_ranks(total_ms, total_e2e, wall_ms, wall_e2e)
    value = world * RAYS * K / (total_ms * 1e-3)
    e2e = world * RAYS * K / (total_e2e * 1e-3)
    time.sleep(0.05)
    clk.__exit__(None, None, None)

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            rps, secs = cpu_train_step_rays_per_s(RAYS, 5, warmup=1)
            cpu = {"value": rps, "unit": "rays/s", "cores": os.cpu_count(), "kind": "port",
                   "sample": f"oracle G1 train step (fwd+bwd+Adam) on the full batch, {RAYS} rays x {SAMPLES} samples, median of 5 ({secs:.1f} s)"}
        tensor_peak = peaks["tensor_sustained"]
        line = {
            "metric": "rays/sec (train fwd+bwd)", "value": value, "unit": "rays/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32" if args.precision == "fp32" else "bf16", "data": "synthetic",
            "config": bench_config(world),
            "e2e": {"value": e2e, "unit": "rays/s", "h2d_bytes_per_step": RAYS * 9 * 4, "d2h_bytes_per_step": 4,
                    "ms_per_step": total_e2e / K, "wall_ms_per_step": wall_e2e / K,
                    "what": "Trainer.step on a batch copied from pinned host memory EVERY step and the loss copied back EVERY step, through "
                            "train.HostBatchPipeline: the copy of batch i+1 is issued on a copy stream before step i's kernels, the "
                            "loss read-back on a second stream; all K copies and K read-backs are inside the timed bracket"},
            "wall_ms_per_step": wall_ms / K,
            "gpu_launches": launches_per_step * K,
            "clocks": clk.summary(),
            "roofline": roofline,
            "roofline_mlp_step": {"bound": "tensor", "kernel": mlp_kernel_name, "achieved": mlp_tflops, "peak": tensor_peak, "unit": "TFLOP/s",
                                  "frac": mlp_tflops / tensor_peak, "ms": ms_mlp,
                                  "peak_source": peaks["src"] + " bf16 sustained (kernels back to back inside a step)"},
            "roofline_hbm_design": roofline_hbm_design,
            "roofline_hbm": {"composite_fwd": {"achieved": comp_fwd_gbs, "frac": comp_fwd_gbs / peaks["hbm"], "ms": ms_cf},
                             "composite_bwd": {"achieved": comp_bwd_gbs, "frac": comp_bwd_gbs / peaks["hbm"], "ms": ms_cb},
                             "sample_host_u": {"achieved": samp_u_gbs, "frac": samp_u_gbs / peaks["hbm"], "ms": ms_su},
                             "sample_philox": {"achieved": samp_p_gbs, "frac": samp_p_gbs / peaks["hbm"], "ms": ms_sp,
                                               "note": "in-kernel Philox4x32-7: 14 IMAD.WIDE per 4 samples on the fma pipe (0.50 of peak with 10 rounds in round 1)"},
                             "peak": peaks["hbm"], "unit": "GB/s", "rays": cb, "samples": s},
            "fp32" if args.precision == "bf16" else "bf16": other_prec,
            "fp32_grade_train": x3_train,
            "api_loop": api_loop,
            "render": {"value": render_rps, "unit": "rays/s", "workload": f"{rh}x{rw} x {SAMPLES} samples render_image, one GPU",
                       "mlp_tflops": render_tflops, "frac_tensor": render_tflops / tensor_peak,
                       "frac_tensor_burst": render_tflops / peaks["tensor_burst"], "tensor_pipe_pct_ncu": rncu.get("tensor_pipe_pct"),
                       "ncu_source": rncu.get("source")},
            "render_fp32_grade": render_fp32_grade,
            "render_sharded": render_sharded,
            "render_d3": d3,
            "train_c4": c4["train"], "render_c4": c4["render"],
            "dp_check": dp_check, "allreduce": allreduce_us,
            "exchange": (trainer.exchange.describe() if trainer.exchange is not None else
                         ("ncclAllReduce(flat gradient) + adam_kernel" if world > 1 else "none (one GPU): adam_kernel")),
            "kernels": kernels,
            "loss": loss_val,
        }
        if isinstance(head, dict):
            line.update(head if "error" not in head else {"head": head})
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="bf16", choices=["fp32", "bf16"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--all-sections", action="store_true", help="also run the HEAD sections at N > 1")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
