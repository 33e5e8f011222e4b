"""Multi-GPU plumbing: rays shard with no communication; training all-reduces the
flat gradient buffer once per step (done inside `Trainer.step`)."""
import os

import torch
import torch.distributed as dist

from ._lib import check, lib, stream


def shard_range(total, rank, world_size):
    """Contiguous slice [begin, end) of `total` flattened ray ids owned by `rank`."""
    per = (total + world_size - 1) // world_size
    begin = min(rank * per, total)
    return begin, min(begin + per, total)


def render_image_sharded(model, camera_o, camera_r, image_h, image_w, focal_length, near, far, num_samples,
                         gather=True, group=None, **kw):
    """Every rank renders its slice of the flattened ray ids; no collective is needed for the
    render itself.  gather=True assembles the frame on all ranks with one all_gather."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    total = camera_o.shape[0] * image_h * image_w
    begin, end = shard_range(total, rank, world)
    part = model.render_image(camera_o, camera_r, image_h, image_w, focal_length, near, far, num_samples,
                              ray_begin=begin, ray_end=end, **kw)
    if not gather or world == 1:
        return part if world > 1 else part.reshape(camera_o.shape[0], image_h, image_w, 3)
    per = (total + world - 1) // world
    buf = torch.zeros(per, 3, device=part.device, dtype=part.dtype)
    buf[:part.shape[0]] = part
    out = torch.empty(world * per, 3, device=part.device, dtype=part.dtype)
    dist.all_gather_into_tensor(out, buf, group=group)
    return out[:total].reshape(camera_o.shape[0], image_h, image_w, 3)


def broadcast_parameters(model, src=0, group=None):
    """Replicate the flat parameter vector from `src` (one broadcast instead of 44)."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.broadcast(model.flat_parameters(), src=src, group=group)


class FusedExchange:
    """The training exchange step as ONE kernel over NVLink/NVSwitch peer memory (csrc/dp_fused.cu): barrier -> all-reduce of
    the flat gradient -> Adam.  Two gradient buffers in symmetric memory (`torch.distributed._symmetric_memory`: the same
    allocation on every rank, peer-mapped, multicast-bound where the fabric offers NVLS) are used alternately, and a
    symmetric flag block carries the per-step epoch barrier.  `FusedExchange.create` returns None when symmetric memory
    cannot be set up on this system (the Trainer then keeps the NCCL all-reduce + Adam launch); NERF_B200_DP=nccl forces
    that, NERF_B200_DP=p2p keeps the fused kernel but reads the peers with plain loads instead of multimem."""

    def __init__(self, n, grads, gh, flags, fh, group):
        self.n, self.n_pad = n, grads.numel() // 2
        self.grads, self.gh, self.flags, self.fh, self.group = grads, gh, flags, fh, group
        self.rank, self.world = gh.rank, gh.world_size
        mode = os.environ.get("NERF_B200_DP", "")
        mc = int(getattr(gh, "multicast_ptr", 0) or 0)
        self.multicast_ptr = 0 if mode == "p2p" else mc
        self.epoch = 0
        # the handle's pointers address the start of the symmetric ALLOCATION; the tensor may sit at an offset inside it
        self.g_off = self._offset_elems(gh, grads, 4)
        self.f_off = self._offset_elems(fh, flags, 8)

    @staticmethod
    def _offset_elems(h, t, itemsize):
        base = int(h.buffer_ptrs[h.rank])
        off = t.data_ptr() - base
        if off < 0 or off % 16 or off + t.numel() * itemsize > int(h.buffer_size):
            raise RuntimeError("symmetric-memory handle does not cover the tensor")
        return off // itemsize

    @classmethod
    def create(cls, n, device, group=None):
        if os.environ.get("NERF_B200_DP", "") == "nccl" or n % 4:
            return None
        try:
            import torch.distributed._symmetric_memory as symm_mem
            grp = group if group is not None else dist.group.WORLD
            n_pad = (n + 1023) // 1024 * 1024
            grads = symm_mem.empty(2 * n_pad, dtype=torch.float32, device=device)
            gh = symm_mem.rendezvous(grads, grp)
            flags = symm_mem.empty(64, dtype=torch.int64, device=device)
            fh = symm_mem.rendezvous(flags, grp)
            grads.zero_(); flags.zero_()
            torch.cuda.synchronize(device)
            dist.barrier(group=group)                           # every rank's flags are zero before anybody publishes
            ok = torch.ones(1, device=device)
        except Exception as e:                                  # noqa: BLE001 -- no symmetric memory here: say so once, use NCCL
            if dist.get_rank(group) == 0:
                print(f"nerf_b200: fused NVLink exchange unavailable ({type(e).__name__}: {e}); using ncclAllReduce + Adam")
            ok, grads = torch.zeros(1, device=device), None
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)  # all ranks or none
        if ok.item() == 0 or grads is None:
            return None
        return cls(n, grads, gh, flags, fh, group)

    def buffer(self):
        """The gradient buffer the NEXT exchange reduces (write the step's gradient here)."""
        h = (self.epoch + 1) & 1
        return self.grads[h * self.n_pad:h * self.n_pad + self.n]

    def allreduce_adam(self, params, exp_avg, exp_avg_sq, step, lr, betas, eps):
        self.epoch += 1
        h = self.epoch & 1
        check(lib().nerf_dp_allreduce_adam(self.multicast_ptr or None, self.gh.buffer_ptrs_dev, self.fh.buffer_ptrs_dev,
                                           self.g_off + h * self.n_pad, self.f_off, self.rank, self.world, self.epoch,
                                           params.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(), self.n, step,
                                           lr, betas[0], betas[1], eps, 1.0 / self.world, stream()), "nerf_dp_allreduce_adam")

    def describe(self):
        return ("one kernel: epoch barrier + " + ("NVLS multimem.ld_reduce" if self.multicast_ptr else "one-shot P2P loads")
                + " + Adam over symmetric memory")
