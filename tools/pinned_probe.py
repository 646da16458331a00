"""PCIe copy time to/from pinned buffers as a function of allocation order (pinned vs device allocations)."""
import time, torch
torch.cuda.init()
d = torch.empty(19_198_080 // 4 * 4 // 4, device="cuda")   # 19.2 MB
def t_copy(dst, src, n=20):
    dst.copy_(src, non_blocking=True); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
a = torch.empty_like(d, device="cpu").pin_memory()
print("D2H -> a (fresh pinned)              %.3f ms" % t_copy(a, d))
print("H2D <- a                             %.3f ms" % t_copy(d, a))
big = torch.empty(60_000_000, device="cuda")
print("D2H -> a after a cudaMalloc          %.3f ms" % t_copy(a, d))
b = torch.empty_like(d, device="cpu").pin_memory()
print("D2H -> b (new pinned)                %.3f ms" % t_copy(b, d))
print("D2H -> a (older pinned)              %.3f ms" % t_copy(a, d))
big2 = torch.empty(70_000_000, device="cuda")
print("D2H -> b after another cudaMalloc    %.3f ms" % t_copy(b, d))
del big, big2; torch.cuda.empty_cache()
print("D2H -> b after cudaFree              %.3f ms" % t_copy(b, d))
c = torch.empty_like(d, device="cpu").pin_memory()
print("D2H -> c (new pinned)                %.3f ms" % t_copy(c, d))
print("D2H -> b                             %.3f ms" % t_copy(b, d))
time.sleep(0.5)
print("D2H -> c after 0.5 s                 %.3f ms" % t_copy(c, d))
