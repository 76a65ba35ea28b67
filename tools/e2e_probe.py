"""Under torchrun: where does the multi-GPU end-to-end rate go?  Per rank, all ranks at once:
(a) H2D of the frame batch with a concurrent D2H of the result size, (b) orbfe_extract_batch from pinned host buffers."""
import os, sys, time, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
import orb_slam2_annotate_b200 as orbfe
from orb_slam2_annotate_b200 import synth
def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
B, W, H = 1024, 640, 480
h = torch.empty(B * W * H, dtype=torch.uint8).pin_memory(); d = torch.empty_like(h, device="cuda")
ho = torch.empty(70_000_000, dtype=torch.uint8).pin_memory(); do = torch.empty_like(ho, device="cuda")
sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
def both():
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(sa)
    with torch.cuda.stream(sa):
        d.copy_(h, non_blocking=True)
    with torch.cuda.stream(sb):
        ho.copy_(do, non_blocking=True)
    e1.record(sa); torch.cuda.synchronize()
    return h.numel() / (e0.elapsed_time(e1) * 1e-3) / 1e9
both(); barrier()
g = max(both() for _ in range(3))
frames = np.stack([synth.frame(W, H, s) for s in range(8)])
hf = orbfe.pinned_empty((B, H, W), np.uint8)
for i in range(B):
    hf[i] = frames[i % 8]
ex = orbfe.ORBextractor(1000)
cap = ex.max_keypoints()
out = (orbfe.pinned_empty((B, cap), orbfe.KP_DTYPE), orbfe.pinned_empty((B, cap, 32), np.uint8), orbfe.pinned_empty((B,), np.int32))
ex.extract_batch(hf, out=out); ex.extract_batch(hf, out=out)
barrier()
t0 = time.perf_counter()
for _ in range(5):
    ex.extract_batch(hf, out=out)
dt = (time.perf_counter() - t0) / 5
barrier()
print(f"rank {rank}/{world}: H2D with concurrent D2H {g:.1f} GB/s; extract_batch {dt*1e3:.2f} ms/step = {B/dt:.0f} frames/s", flush=True)
if world > 1:
    dist.destroy_process_group()
