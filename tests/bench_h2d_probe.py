"""H2D bandwidth probe for the host-buffer entry point (not a pytest file): one cudaMemcpyAsync, a 2-D copy with equal
pitches, and the same bytes split over 2 / 4 streams, from pinned memory."""
import time
import torch
n = 4 << 30
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
def timed(fn, reps=3):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return n * reps / (time.perf_counter() - t0) / 1e9
print("1 stream, 1-D      %.1f GB/s" % timed(lambda: d.copy_(h, non_blocking=True)))
h2, d2 = h.view(32768, -1), d.view(32768, -1)
print("1 stream, 2-D view %.1f GB/s" % timed(lambda: d2.copy_(h2, non_blocking=True)))
for k in (2, 4, 8):
    streams = [torch.cuda.Stream() for _ in range(k)]
    hs, ds = h.chunk(k), d.chunk(k)
    def go():
        for s, a, b in zip(streams, hs, ds):
            with torch.cuda.stream(s): b.copy_(a, non_blocking=True)
    print("%d streams          %.1f GB/s" % (k, timed(go)))
import subprocess
print(subprocess.run(["nvidia-smi", "--query-gpu=pcie.link.gen.current,pcie.link.width.current,pcie.link.gen.max", "--format=csv"], capture_output=True, text=True).stdout)
print(subprocess.run(["bash", "-c", "lscpu | grep -E 'NUMA|Model name|Socket'; numactl -H 2>/dev/null | head -5"], capture_output=True, text=True).stdout)
