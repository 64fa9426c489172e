"""Does a pinned H2D copy on another stream slow concurrent kernels on this box?
Cases: compute-bound (fp32 matmul), HBM-bound (big elementwise), L2-resident elementwise."""
import torch, time
dev = torch.device("cuda", 0)
host = torch.empty(64 * 1024 * 1024 // 4, dtype=torch.float32).pin_memory()
dst = torch.empty_like(host, device=dev)
cs = torch.cuda.Stream()
a = torch.randn(4096, 4096, device=dev); b = torch.randn(4096, 4096, device=dev)
big = torch.randn(256 * 1024 * 1024 // 4, device=dev); small = torch.randn(8 * 1024 * 1024 // 4, device=dev)

def bench(fn, copy, n=30):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for i in range(n):
        if copy:
            with torch.cuda.stream(cs):
                dst.copy_(host, non_blocking=True)
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

for name, fn in (("matmul fp32 4096", lambda: torch.mm(a, b)), ("elementwise 256MB", lambda: big.mul_(1.0001)),
                 ("elementwise 8MB x20", lambda: [small.mul_(1.0001) for _ in range(20)])):
    print(f"{name:24s} alone {bench(fn, False):7.3f} ms   with H2D {bench(fn, True):7.3f} ms")
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record(cs)
with torch.cuda.stream(cs):
    for _ in range(10): dst.copy_(host, non_blocking=True)
e1.record(cs); torch.cuda.synchronize()
print("H2D 64 MiB alone", e0.elapsed_time(e1) / 10, "ms")
