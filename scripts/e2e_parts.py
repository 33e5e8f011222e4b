"""Where the end-to-end step differs from the device-resident one: the same 200 steps with the pipeline pieces added one at a
time (CUDA events over the region + the host time spent issuing it).  The GPU power-caps over the first seconds of load, so
compare a line with its neighbours, not with the first one."""
import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import nerf_pth_state_dict, synthetic_batch
from nerf_b200.model import NeRF
from nerf_b200.train import Trainer, HostBatchPipeline
b, s = 4096, 64
m = NeRF(normalize_position=6.0, precision="bf16").cuda(); m.load_state_dict(nerf_pth_state_dict())
tr = Trainer(m, num_samples=s, density_noise_std=1.0, randomly_sample=True, seed=1)
pool_h = []
for i in range(8):
    o, d = synthetic_batch(b, i); tgt = torch.rand(b, 3)
    pool_h.append(torch.stack([o, d, tgt]).pin_memory())
pool_d = [tuple(x.cuda()) for x in pool_h]
K = 200
def region(f):
    for i in range(10): f(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.perf_counter(); e0.record()
    for i in range(K): f(i)
    cpu = time.perf_counter() - w0
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / K, cpu / K * 1e3
def resident(i):
    tr.step(*pool_d[i % 8])
ev = [torch.cuda.Event() for _ in range(4)]
def resident_events(i):
    cur = torch.cuda.current_stream()
    cur.wait_event(ev[0]) if i else None
    tr.step(*pool_d[i % 8])
    ev[0].record(cur); ev[1].record(cur)
pipe = HostBatchPipeline(torch.device("cuda"), like=pool_h[0])
state = {"first": True}
def e2e(i, loss=True, copy=True):
    if state["first"]:
        pipe.submit(pool_h[0]); state["first"] = False
    pipe.submit(pool_h[(i + 1) % 8])
    o, d, t = pipe.get()
    l = tr.step(o, d, t)
    pipe.done(l if loss else None, ring=tr.loss_ring)
h2d = torch.cuda.Stream(); dbuf = torch.empty_like(pool_d[0][0].new_empty(3, b, 3))
def copy_only(i):
    with torch.cuda.stream(h2d):
        dbuf.copy_(pool_h[i % 8], non_blocking=True)
    tr.step(*pool_d[i % 8])
for name, f in (("resident", resident), ("resident + 1 wait + 2 records", resident_events), ("resident + unsynchronised h2d copy on a side stream", copy_only),
                ("e2e, no loss read-back", lambda i: e2e(i, loss=False)), ("e2e", e2e), ("resident again", resident)):
    g, c = region(f)
    print(f"{name:55s} device {g:.4f} ms/step   host issue {c:.4f} ms/step")
