"""Exploration: raw PCIe copy rates on the box (one direction, both directions at once) vs the e2e pipeline."""
import torch, time
torch.cuda.set_device(0)
n = 1 << 30
h_in = torch.empty(n, dtype=torch.uint8).pin_memory(); h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda"); d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(h2d, d2h, reps=5):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize(); return reps * n / (time.perf_counter() - t0) / 1e9
run(True, True, 1)
print("H2D alone %.1f GB/s" % run(True, False))
print("D2H alone %.1f GB/s" % run(False, True))
print("both at once: %.1f GB/s each way" % run(True, True))
for chunk_mb in (16, 64, 128, 512):
    c = chunk_mb << 20
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for r in range(3):
        for o in range(0, n, c):
            with torch.cuda.stream(s1): d_in[o:o + c].copy_(h_in[o:o + c], non_blocking=True)
            with torch.cuda.stream(s2): h_out[o:o + c].copy_(d_out[o:o + c], non_blocking=True)
    torch.cuda.synchronize()
    print("both at once in %d MB chunks: %.1f GB/s each way" % (chunk_mb, 3 * n / (time.perf_counter() - t0) / 1e9))
