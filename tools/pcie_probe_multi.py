"""Concurrent H2D probe under torchrun: every rank copies the bench's 315 MB frame batch from pinned memory at the
same time; prints per-rank and aggregate GB/s plus where the GPU and the process sit (NUMA node, CPU affinity)."""
import os, time, torch, torch.distributed as dist
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl")
def numa_of_gpu(i):
    try:
        bus = torch.cuda.get_device_properties(i).pci_bus_id
        dom = torch.cuda.get_device_properties(i).pci_domain_id
        dev = torch.cuda.get_device_properties(i).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/numa_node"
        return int(open(path).read())
    except Exception as e:
        return f"? ({e})"
aff = sorted(os.sched_getaffinity(0))
nodes = {}
try:
    for n in os.listdir("/sys/devices/system/node"):
        if n.startswith("node"):
            nodes[n] = open(f"/sys/devices/system/node/{n}/cpulist").read().strip()
except Exception:
    pass
h = torch.empty(1024 * 480 * 640, dtype=torch.uint8).pin_memory()
d = torch.empty_like(h, device="cuda")
d.copy_(h, non_blocking=True); torch.cuda.synchronize()
def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
res = []
for trial in range(3):
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(4):
        d.copy_(h, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    res.append(4 * h.numel() / (e0.elapsed_time(e1) * 1e-3) / 1e9)
g = max(res)
t = torch.tensor([g], device="cuda")
if world > 1:
    dist.all_reduce(t)
print(f"rank {rank}: gpu numa {numa_of_gpu(lr)} affinity {aff[:4]}..{aff[-1]} ({len(aff)} cpus) H2D {g:.1f} GB/s", flush=True)
barrier()
if rank == 0:
    print(f"world {world}: aggregate H2D {t.item():.1f} GB/s; numa nodes {nodes}", flush=True)
if world > 1:
    dist.destroy_process_group()
