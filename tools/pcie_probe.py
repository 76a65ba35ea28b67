"""PCIe probe: H2D of the bench's frame batch in chunks, alone and with a concurrent D2H stream of the result size."""
import sys, torch
B, W, H = 1024, 640, 480
h = torch.empty((B, H, W), dtype=torch.uint8).pin_memory()
d = torch.empty((B, H, W), dtype=torch.uint8, device="cuda")
ho = torch.empty(70_000_000, dtype=torch.uint8).pin_memory()
do = torch.empty(70_000_000, dtype=torch.uint8, device="cuda")
sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
def run(chunk, with_d2h):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = B // chunk
    oc = ho.numel() // n
    e0.record(sa)
    for c in range(n):
        with torch.cuda.stream(sa):
            d[c * chunk:(c + 1) * chunk].copy_(h[c * chunk:(c + 1) * chunk], non_blocking=True)
        if with_d2h:
            with torch.cuda.stream(sb):
                ho[c * oc:(c + 1) * oc].copy_(do[c * oc:(c + 1) * oc], non_blocking=True)
    e1.record(sa)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)
for chunk in (1024, 256, 64, 16):
    for w in (False, True):
        run(chunk, w)
        t = min(run(chunk, w) for _ in range(5))
        print(f"chunk {chunk:5d} d2h {int(w)}: H2D {t:.3f} ms = {B*W*H/t/1e6:.1f} GB/s")
