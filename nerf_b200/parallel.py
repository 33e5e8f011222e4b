"""Multi-GPU plumbing: rays shard with no communication; training all-reduces the
flat gradient buffer once per step (done inside `Trainer.step`)."""
import torch
import torch.distributed as dist


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
