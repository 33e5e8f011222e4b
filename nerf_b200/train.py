"""The fused training step (reference loop body: train_conditional_nerf.py:120-135,
example.ipynb cell 8 lines 10-23) without autograd: sample -> encode -> MLP ->
noise -> composite -> MSE -> composite backward -> MLP backward -> [all-reduce] -> Adam.
Every arithmetic stage is one of our kernels; torch only owns memory, streams and
(multi-GPU) the NCCL all-reduce of the flat gradient buffer."""
import torch
import torch.distributed as dist

from . import ops
from ._lib import check, lib, ptr, stream


class HostBatchPipeline:
    """Host -> device input pipeline for the training loop: batches arrive in PINNED host memory and are copied into
    one of `depth` device buffers on a copy stream while the previous step computes (what a DataLoader with
    pin_memory + non_blocking copies does for the reference's loop, train_conditional_nerf.py:100-104); the loss of
    each step is read back into pinned memory on a second stream.  Every step still moves its own batch and its own
    loss across PCIe; only the waiting overlaps.

        pipe = HostBatchPipeline(device, like=host_batch)
        pipe.submit(host_batches[0])
        for i in range(n):
            if i + 1 < n: pipe.submit(host_batches[i + 1])     # copy of the next batch runs under this step
            o, d, target = pipe.get()                          # current stream waits for this batch's copy only
            loss = trainer.step(o, d, target)
            pipe.done(loss, ring=trainer.loss_ring)            # frees the buffer; schedules the loss read-back
        pipe.losses()                                          # pinned host tensor with the losses read back so far

    `ring` tells the pipeline when the producer overwrites the storage of `loss`: `ring` steps later (default 1: the very next
    step, e.g. a single device scalar).  The step that overwrites it is made to wait for its read-back; with ring = 1 that is
    an event round trip and a 4-byte copy between EVERY two steps (~55 us of a 1.5 ms step, measured), with the Trainer's ring
    of 4 slots the read-back waited for is three steps old and long complete.
    """

    def __init__(self, device, like, depth=2, max_losses=1 << 16):
        self.bufs = [torch.empty(like.shape, dtype=like.dtype, device=device) for _ in range(depth)]
        self.copied = [torch.cuda.Event() for _ in range(depth)]
        self.freed = [torch.cuda.Event() for _ in range(depth)]
        self.h2d, self.d2h = torch.cuda.Stream(device), torch.cuda.Stream(device)
        self.host_losses = torch.zeros(max_losses).pin_memory()
        self.stepped, self.loss_read = torch.cuda.Event(), torch.cuda.Event()
        self.read_events = [torch.cuda.Event() for _ in range(8)]
        self.reads = []                                            # (step at which the loss storage is overwritten, read-back event)
        self.head = self.tail = self.n_losses = 0
        self.pending = [False] * depth

    def submit(self, host_batch):
        k = self.head % len(self.bufs)
        if self.pending[k]:
            raise RuntimeError("HostBatchPipeline: more batches in flight than buffers")
        with torch.cuda.stream(self.h2d):
            if self.head >= len(self.bufs):
                self.h2d.wait_event(self.freed[k])             # the step that read this buffer has been issued and finished
            self.bufs[k].copy_(host_batch, non_blocking=True)
            self.copied[k].record(self.h2d)
        self.pending[k] = True
        self.head += 1

    def get(self):
        k = self.tail % len(self.bufs)
        if not self.pending[k]:
            raise RuntimeError("HostBatchPipeline: get() without a submitted batch")
        cur = torch.cuda.current_stream()
        cur.wait_event(self.copied[k])
        keep = []
        for reuse, ev in self.reads:                           # read-backs whose source the coming step overwrites precede it
            if reuse <= self.n_losses:
                if not ev.query():                             # (a wait on a finished event is still a stream operation between
                    cur.wait_event(ev)                         #  two kernels; an old read-back has long finished)
            else:
                keep.append((reuse, ev))
        self.reads = keep
        return tuple(self.bufs[k])

    def done(self, loss=None, ring=1):
        if not 1 <= ring < len(self.read_events):
            raise ValueError("HostBatchPipeline.done: ring must be in 1..7")
        k = self.tail % len(self.bufs)
        cur = torch.cuda.current_stream()
        self.freed[k].record(cur)                              # one record per step: "the buffer is free" = "the loss is written"
        self.stepped = self.freed[k]
        self.pending[k] = False
        self.tail += 1
        if loss is not None:
            with torch.cuda.stream(self.d2h):
                self.d2h.wait_event(self.stepped)
                self.host_losses[self.n_losses % self.host_losses.numel():][:1].copy_(loss.detach().reshape(1), non_blocking=True)
                self.loss_read = self.read_events[self.n_losses % len(self.read_events)]
                self.loss_read.record(self.d2h)
            self.reads.append((self.n_losses + ring, self.loss_read))
            self.n_losses += 1

    def losses(self):
        self.d2h.synchronize()
        return self.host_losses[:min(self.n_losses, self.host_losses.numel())]


class Trainer:
    LOSS_RING = 4      # loss slots of the tensor-core step, used in turn: step i's loss is overwritten by step i + 4

    def __init__(self, model, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, near=2.0, far=6.0,
                 num_samples=64, density_noise_std=1.0, randomly_sample=True, seed=0,
                 precision=None, process_group=None, fused_exchange=True):
        self.model = model
        self.lr, self.betas, self.eps = lr, betas, eps
        self.near, self.far, self.num_samples = near, far, num_samples
        self.noise_std, self.jitter, self.seed = density_noise_std, randomly_sample, seed
        self.precision = ops.PRECISIONS[precision or model.precision]
        # bf16x3 (fp32-grade operands on tcgen05) trains the unconditional model; conditional models of that mode step in fp32
        if self.precision == ops.PREC_BF16X3 and (model.density_inputs != 3 or model.color_inputs != 3):
            self.precision = ops.PREC_FP32
        self.group = process_group
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank(process_group) if self.world > 1 else 0
        flat = model.flat_parameters()
        # data-parallel exchange: ONE kernel (barrier + all-reduce over NVLink peer memory + Adam, parallel.FusedExchange)
        # when symmetric memory is available, else ncclAllReduce of the flat gradient followed by the Adam launch
        self.exchange = None
        if self.world > 1 and fused_exchange:
            from .parallel import FusedExchange
            self.exchange = FusedExchange.create(flat.numel(), flat.device, process_group)
        self.flat_grad = self.exchange.buffer() if self.exchange else torch.zeros_like(flat)
        self.exp_avg = torch.zeros_like(flat)
        self.exp_avg_sq = torch.zeros_like(flat)
        self.step_count = 0
        self.rays_seen = 0
        # how many steps later the storage of the loss the LAST step returned is overwritten (HostBatchPipeline.done(loss,
        # ring=...)): the fp32 step returns a fresh tensor whose block the allocator may hand out again at once
        self.loss_ring = 1
        self._loss_k = 0
        self._ws = None
        self._tc = None                 # bf16 path: buffers keyed by (B, S)
        self.kernel_launches = 0        # kernels launched by the last step() (our .so only)

    def _workspace(self, p):
        dx, dd = self.model.density_inputs, self.model.color_inputs
        n = lib().nerf_g1_mlp_workspace_bytes(p, dx, dd, 1)
        if self._ws is None or self._ws.numel() < n:
            self._ws = torch.empty(n, dtype=torch.uint8, device=self.model.flat_parameters().device)
        return self._ws, n

    # ------------------------------------------------------------ bf16 (tcgen05) step ------
    def _tc_buffers(self, b, s, dev):
        dx = self.model.density_inputs
        x3 = self.precision == ops.PREC_BF16X3
        if self._tc is None or self._tc["key"] != (b, s, x3):
            L = lib()
            n = L.nerf_g1_train_x3_workspace_bytes(b, s) if x3 else L.nerf_g1_train_workspace_bytes(b, s, dx)
            e = lambda *shape: torch.empty(*shape, device=dev, dtype=torch.float32)
            self._tc = {
                "key": (b, s, x3), "n": n, "ws": torch.empty(n, dtype=torch.uint8, device=dev),
                "packed": torch.empty(L.nerf_g1_packed_x3_bytes() if x3 else L.nerf_g1_packed_bytes(dx, 3), dtype=torch.uint8, device=dev),
                "packed_bwd": torch.empty(L.nerf_g1_packed_bwd_x3_bytes() if x3 else L.nerf_g1_packed_bwd_bytes(dx, 3),
                                          dtype=torch.uint8, device=dev),
                "t": e(b, s), "sigma": e(b, s), "rgb": e(b, s, 3), "d_sigma": e(b, s), "d_rgb": e(b, s, 3),
                "pixels": e(b, 3), "g_pixels": e(b, 3), "loss": e(self.LOSS_RING),
            }
        return self._tc

    def _step_bf16(self, rays_o, rays_d, target, u, noise, apply_update, states_x=None):
        """sample+encode+MLP forward (one tcgen05 kernel, activations kept as bf16 slab images) ->
        noise -> composite -> MSE -> composite backward -> tcgen05 data-gradient kernel ->
        tcgen05 weight-gradient GEMM -> unfold -> [all-reduce] -> Adam.  fp32 master weights."""
        m = self.model
        flat = m.flat_parameters()
        b, s = rays_o.shape[0], self.num_samples
        dev = flat.device
        L = lib()
        z = self._tc_buffers(b, s, dev)
        loss = z["loss"][self._loss_k:self._loss_k + 1]
        self._loss_k = (self._loss_k + 1) % self.LOSS_RING
        self.loss_ring = self.LOSS_RING
        st = stream()
        ray_base = self.rays_seen * self.world + self.rank * b
        dx = m.density_inputs
        sx = ops._f32c(states_x) if dx > 3 else None
        x3 = self.precision == ops.PREC_BF16X3
        ladder = m._ladder(self.near, self.far, s, dev)
        if x3:
            check(L.nerf_g1_pack_weights_x3(ptr(flat), ptr(z["packed"]), st), "nerf_g1_pack_weights_x3")
            check(L.nerf_g1_pack_weights_bwd_x3(ptr(flat), ptr(z["packed_bwd"]), st), "nerf_g1_pack_weights_bwd_x3")
            check(L.nerf_g1_train_fwd_x3(ptr(z["packed"]), ptr(rays_o), ptr(rays_d), b, s, self.near, self.far, ptr(ladder), None, ptr(u),
                                         int(self.jitter and u is None), self.seed, ray_base,
                                         float(m.normalize_position), int(m.normalize_direction),
                                         ptr(z["t"]), ptr(z["sigma"]), ptr(z["rgb"]), ptr(z["ws"]), z["n"], st), "nerf_g1_train_fwd_x3")
        else:
            check(L.nerf_g1_pack_weights(ptr(flat), dx, 3, ptr(z["packed"]), st), "nerf_g1_pack_weights")
            check(L.nerf_g1_pack_weights_bwd(ptr(flat), dx, 3, ptr(z["packed_bwd"]), st), "nerf_g1_pack_weights_bwd")
            check(L.nerf_g1_train_fwd_tc(ptr(z["packed"]), dx, 3, ptr(rays_o), ptr(rays_d), b, s, self.near, self.far,
                                         ptr(ladder), None, ptr(u),
                                         int(self.jitter and u is None), self.seed, ray_base,
                                         float(m.normalize_position), int(m.normalize_direction), ptr(sx),
                                         ptr(z["t"]), ptr(z["sigma"]), ptr(z["rgb"]), ptr(z["ws"]), z["n"], st),
                  "nerf_g1_train_fwd_tc")
        launches = 4 + (5 if dx > 3 else 0)     # conditional: state weights, PE(states), ray bias; dz ray sums, state gradients
        std = float(self.noise_std)
        if noise is not None:                   # host-supplied draw (parity mode)
            z["sigma"] += noise.view(b, s); std = 0.0
        if s <= 256:                            # noise + composite + MSE + composite backward: one launch
            check(L.nerf_composite_mse_train(ptr(z["sigma"]), ptr(z["rgb"]), ptr(rays_o), ptr(rays_d), ptr(z["t"]), ptr(target),
                                             b, s, std, self.seed, ray_base, ptr(z["pixels"]), ptr(loss),
                                             ptr(z["d_sigma"]), ptr(z["d_rgb"]), st), "nerf_composite_mse_train")
            launches += 1
        else:
            if std != 0.0:
                ops.add_density_noise_(z["sigma"], std, self.seed, ray_base); launches += 1
            check(L.nerf_composite_fwd(ptr(z["sigma"]), ptr(z["rgb"]), None, ptr(rays_o), ptr(rays_d), ptr(z["t"]), b, s,
                                       ptr(z["pixels"]), None, None, None, st), "nerf_composite_fwd")
            check(L.nerf_mse_loss(ptr(z["pixels"]), ptr(target), b * 3, ptr(loss), ptr(z["g_pixels"]), st), "nerf_mse_loss")
            check(L.nerf_composite_bwd(ptr(z["sigma"]), ptr(z["rgb"]), None, ptr(rays_o), ptr(rays_d), ptr(z["t"]), b, s,
                                       ptr(z["g_pixels"]), None, None, None, ptr(z["pixels"]), None, None,
                                       ptr(z["d_sigma"]), ptr(z["d_rgb"]), st),
                  "nerf_composite_bwd")
            launches += 3
        if x3:
            check(L.nerf_g1_train_bwd_x3(ptr(z["packed"]), ptr(z["packed_bwd"]), ptr(flat), ptr(z["d_sigma"]), ptr(z["d_rgb"]), b, s,
                                         ptr(self.flat_grad), ptr(z["ws"]), z["n"], 0, st), "nerf_g1_train_bwd_x3")
            launches += 5                                           # bwd, 3 x wgrad, unfold
        else:
            check(L.nerf_g1_train_bwd_tc(ptr(z["packed"]), ptr(z["packed_bwd"]), ptr(flat), dx, 3, ptr(z["d_sigma"]),
                                         ptr(z["d_rgb"]), b, s, ptr(self.flat_grad), ptr(z["ws"]), z["n"], 0, st),
                  "nerf_g1_train_bwd_tc")
            launches += 3                                           # bwd, wgrad, unfold (memset not counted)
        launches += self._exchange_and_update(flat, apply_update)
        self.rays_seen += b
        self.kernel_launches = launches
        self.last_pixels = z["pixels"]
        return loss

    def _exchange_and_update(self, flat, apply_update):
        """[all-reduce of the flat gradient] + Adam.  Returns the number of OUR kernels launched.  With the fused
        exchange `self.flat_grad` keeps this rank's LOCAL gradient (the sum exists only inside the kernel) and the next
        step writes the other symmetric buffer; `apply_update=False` (tests) always takes the NCCL all-reduce."""
        if self.exchange is not None and apply_update and self.world > 1:
            self.step_count += 1
            self.exchange.allreduce_adam(flat, self.exp_avg, self.exp_avg_sq, self.step_count, self.lr, self.betas, self.eps)
            self.last_grad = self.flat_grad
            self.flat_grad = self.exchange.buffer()
            return 1
        if self.world > 1:
            dist.all_reduce(self.flat_grad, group=self.group)        # one NCCL all-reduce of 2.55 MB
        if apply_update:
            self.step_count += 1
            ops.adam_step(flat, self.flat_grad, self.exp_avg, self.exp_avg_sq, self.step_count, self.lr,
                          self.betas, self.eps, grad_scale=1.0 / self.world)
            return 1
        return 0

    def step(self, rays_o, rays_d, target, states_x=None, states_d=None, u=None, noise=None,
             apply_update=True):
        """One optimisation step on a ray batch resident on the device.  Returns the loss
        (1-element device tensor; reading it is the caller's device->host sync).  On the tensor-core paths the tensor is a
        view of one of LOSS_RING slots and is overwritten `self.loss_ring` steps later: `.item()` / `.clone()` it, or hand it
        to HostBatchPipeline.done(loss, ring=self.loss_ring), before then."""
        m = self.model
        if (self.precision in (ops.PREC_BF16, ops.PREC_BF16X3) and states_d is None and m.color_inputs == 3 and self.num_samples >= 32
                and (m.density_inputs == 3 or (states_x is not None and self.precision == ops.PREC_BF16))):
            return self._step_bf16(rays_o.contiguous(), rays_d.contiguous(), target.contiguous(), u, noise, apply_update,
                                   states_x=states_x)
        self.loss_ring = 1
        flat = m.flat_parameters()
        b, s = rays_o.shape[0], self.num_samples
        dx, dd = m.density_inputs, m.color_inputs
        dev = flat.device
        ray_base = self.rays_seen * self.world + self.rank * b      # global ray id for the counter RNG
        launches = 0
        t = ops.sample_linear(self.near, self.far, s, b, dev, u=u, jitter=self.jitter and u is None,
                              seed=self.seed, ray_id_base=ray_base, t_base=m._ladder(self.near, self.far, s, dev))
        launches += 1
        ex, ed, _ = ops.encode_g1(rays_o, rays_d, t, m.normalize_position, states_x, states_d, m.normalize_direction)
        launches += 2
        ws, nbytes = self._workspace(b * s)
        sigma = torch.empty(b, s, device=dev); rgb = torch.empty(b, s, 3, device=dev)
        check(lib().nerf_g1_mlp_fwd(ptr(flat), ptr(ex), ptr(ed), b * s, s, dx, dd, 1,
                                    ptr(sigma), ptr(rgb), ptr(ws), nbytes, stream()), "nerf_g1_mlp_fwd")
        launches += 12
        if noise is not None:
            sigma += noise.view(b, s)
        elif self.noise_std != 0.0:
            ops.add_density_noise_(sigma, self.noise_std, self.seed, ray_base); launches += 1
        pixels = torch.empty(b, 3, device=dev)
        check(lib().nerf_composite_fwd(ptr(sigma), ptr(rgb), None, ptr(rays_o), ptr(rays_d), ptr(t), b, s,
                                       ptr(pixels), None, None, None, stream()), "nerf_composite_fwd")
        loss, g_pixels = ops.mse_loss(pixels, target)
        d_sigma = torch.empty_like(sigma); d_rgb = torch.empty_like(rgb)
        check(lib().nerf_composite_bwd(ptr(sigma), ptr(rgb), None, ptr(rays_o), ptr(rays_d), ptr(t), b, s,
                                       ptr(g_pixels), None, None, None, ptr(pixels), None, None,
                                       ptr(d_sigma), ptr(d_rgb), stream()),
              "nerf_composite_bwd")
        launches += 3
        self.flat_grad.zero_()
        check(lib().nerf_g1_mlp_bwd(ptr(flat), ptr(ex), ptr(ed), b * s, s, dx, dd,
                                    ptr(d_sigma), ptr(d_rgb), ptr(self.flat_grad), ptr(ws), nbytes, stream()),
              "nerf_g1_mlp_bwd")
        launches += 29
        launches += self._exchange_and_update(flat, apply_update)
        self.rays_seen += b
        self.kernel_launches = launches
        self.last_pixels = pixels
        return loss
