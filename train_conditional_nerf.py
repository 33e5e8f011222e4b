"""Conditional (latent-coded) NeRF training entry point, flag-compatible with the reference's
train_conditional_nerf.py (same 13 flags, same files written to --logging-dir), running on the
B200 kernels.  Additions: --precision, --synthetic (the reference's examples/data_for_nerf.pkl is
not redistributable/present), torchrun data parallelism (one rank per GPU; rays
sharded, flat-gradient all-reduce over NCCL)."""
import argparse
import json
import os
import pickle as pkl

import numpy as np
import torch
import torch.distributed as dist

from nerf.dataset import PixelRayDataset
from nerf.model import NeRF
from nerf_b200.parallel import broadcast_parameters
from nerf_b200.train import Trainer


def synthetic_scenes(n_views=24, hw=32, n_states=4, seed=4):
    """Same keys/shapes as the reference pickle: images [N,H,W,3], poses [N,6] = xyz + view
    direction, states [N,Sx] (train_conditional_nerf.py:71-76)."""
    g = np.random.default_rng(seed)
    th = g.uniform(0, 2 * np.pi, n_views); ph = g.uniform(0.2, 1.2, n_views)
    xyz = 4.0 * np.stack([np.cos(th) * np.cos(ph), np.sin(th) * np.cos(ph), np.sin(ph)], -1)
    poses = np.concatenate([xyz, -xyz / np.linalg.norm(xyz, axis=-1, keepdims=True)], -1).astype(np.float32)
    states = g.uniform(-np.pi, np.pi, (n_views, n_states)).astype(np.float32)
    images = g.uniform(0, 1, (n_views, hw, hw, 3)).astype(np.float32)
    return dict(images=images, poses=poses, states=states)


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("--logging-dir", type=str, default="experiment")
    parser.add_argument("--data", type=str, default="examples/data_for_nerf.pkl")
    parser.add_argument("--epochs", type=int, default=100)
    parser.add_argument("--camera-focal-length", type=float, default=50.0)
    parser.add_argument("--camera-ccd-width", type=float, default=36.0)
    parser.add_argument("--batch-size", type=int, default=1024)
    parser.add_argument("--normalize-position", type=float, default=20.0)
    parser.add_argument("--learning-rate", type=float, default=0.0001)
    parser.add_argument("--near-plane", type=float, default=0.0)
    parser.add_argument("--far-plane", type=float, default=20.0)
    parser.add_argument("--num-samples-per-ray", type=int, default=64)
    parser.add_argument("--density-noise-std", type=float, default=1.0)
    parser.add_argument("--log-interval", type=int, default=1000)
    parser.add_argument("--precision", type=str, default="fp32", choices=["fp32", "bf16"])
    parser.add_argument("--synthetic", action="store_true")
    parser.add_argument("--max-iterations", type=int, default=0)
    args = parser.parse_args()

    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        os.makedirs(args.logging_dir, exist_ok=True)
        with open(os.path.join(args.logging_dir, "params.json"), "w") as f:
            json.dump({k: v for k, v in vars(args).items()}, f)

    if args.synthetic or not os.path.exists(args.data):
        data = synthetic_scenes()
    else:
        with open(args.data, "rb") as f:
            data = pkl.load(f)
    images, poses, states_x = data["images"], data["poses"], data["states"]
    H, W = images[0].shape[:2]
    focal_length = float(W) * (args.camera_focal_length / args.camera_ccd_width)
    images = torch.FloatTensor(np.asarray(images)).cuda()
    poses = torch.FloatTensor(np.asarray(poses)).cuda()
    states_x = torch.FloatTensor(np.asarray(states_x)).cuda() / np.pi
    poses = torch.cat([NeRF.direction_to_rotation_matrix(poses[:, 3:]), poses[:, :3].unsqueeze(2)], dim=2)
    test_image, test_pose, test_state_x = images[-1:], poses[-1:], states_x[-1:]
    images, poses, states_x = images[:-1], poses[:-1], states_x[:-1]

    dataset = PixelRayDataset(images, poses, focal_length, states_x=states_x)
    nerf = NeRF(normalize_position=args.normalize_position, density_inputs=3 + states_x.shape[-1],
                precision=args.precision).cuda()
    broadcast_parameters(nerf)
    trainer = Trainer(nerf, lr=args.learning_rate, near=args.near_plane, far=args.far_plane,
                      num_samples=args.num_samples_per_ray, density_noise_std=args.density_noise_std,
                      randomly_sample=True, seed=0)
    psnrs, iternums, rendered_images, ground_truth_images = [], [], [], []
    iteration = -1
    done = False
    for epoch in range(args.epochs):
        gen = torch.Generator().manual_seed(epoch)
        for batch in dataset.batches(args.batch_size, shuffle=True, generator=gen, rank=rank, world_size=world):
            iteration += 1
            loss = trainer.step(batch["rays_o"], batch["rays_d"], batch["pixels"], states_x=batch["states_x"])
            if iteration % args.log_interval == 0 and rank == 0:
                with torch.no_grad():
                    test_render = nerf.render_image(test_pose[..., :3, 3], test_pose[..., :3, :3], H, W, focal_length,
                                                    args.near_plane, args.far_plane, args.num_samples_per_ray,
                                                    states_x=test_state_x)
                psnr = -10.0 * torch.log(((test_render - test_image) ** 2).mean()) / 2.30258509299
                psnrs.append(psnr.cpu().numpy()); iternums.append(iteration)
                rendered_images.append(test_render.cpu().numpy()); ground_truth_images.append(test_image.cpu().numpy())
                print(f"iteration {iteration} loss {loss.item():.5f} psnr {float(psnr):.2f}", flush=True)
                torch.save(nerf.state_dict(), os.path.join(args.logging_dir, "model.pth"))
                np.save(os.path.join(args.logging_dir, "psnrs.npy"), np.asarray(psnrs))
                np.save(os.path.join(args.logging_dir, "iternums.npy"), np.asarray(iternums))
                np.save(os.path.join(args.logging_dir, "rendered_images.npy"), np.asarray(rendered_images))
                np.save(os.path.join(args.logging_dir, "ground_truth_images.npy"), np.asarray(ground_truth_images))
            if args.max_iterations and iteration + 1 >= args.max_iterations:
                done = True
                break
        if done:
            break
    if world > 1:
        dist.destroy_process_group()
