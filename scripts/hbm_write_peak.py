"""Write-only, read-only and copy bandwidth of this GPU with plain torch ops (4 GiB buffers, CUDA events): the denominators for
kernels whose HBM traffic is one-directional (the training forward only writes its stash, the weight-gradient GEMM only reads)."""
import torch
n = 1 << 30                                     # float32 elements = 4 GiB
a = torch.empty(n, dtype=torch.float32, device="cuda"); b = torch.empty_like(a)


def timed(f, it=5):
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it * 1e-3


t = timed(lambda: a.zero_()); print(f"write only (memset): {4 * n / t / 1e12:.2f} TB/s")
t = timed(lambda: a.fill_(1.5)); print(f"write only (fill kernel): {4 * n / t / 1e12:.2f} TB/s")
t = timed(lambda: a.sum()); print(f"read only (sum): {4 * n / t / 1e12:.2f} TB/s")
t = timed(lambda: b.copy_(a)); print(f"copy (read + write bytes): {8 * n / t / 1e12:.2f} TB/s")
