"""How fast does this box move the 1.18 MB fit target host -> device?  One pinned copy, the copy split over 2 / 4 streams, and
write-combined pinned memory (cudaHostAlloc through ctypes).   python tools/h2d_probe.py"""
import ctypes
import time

import torch

n = 768 * 512 * 3
dev = torch.device("cuda:0")
dst = torch.empty(n, dtype=torch.uint8, device=dev)
src = torch.randint(0, 255, (n,), dtype=torch.uint8).pin_memory()
rt = ctypes.CDLL("libcudart.so.12")
p = ctypes.c_void_p()
assert rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(n), ctypes.c_uint(4)) == 0          # cudaHostAllocWriteCombined
wc = torch.frombuffer((ctypes.c_uint8 * n).from_address(p.value), dtype=torch.uint8)
wc.copy_(src)
streams = [torch.cuda.Stream() for _ in range(4)]


def timed(fn, reps=200):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return 1e6 * (time.perf_counter() - t0) / reps


def split(host, k):
    def f():
        c = n // k
        for i in range(k):
            with torch.cuda.stream(streams[i]):
                dst[i * c:(i + 1) * c].copy_(host[i * c:(i + 1) * c], non_blocking=True)
    return f


def raw(host_ptr):
    def f():
        rt.cudaMemcpyAsync(ctypes.c_void_p(dst.data_ptr()), ctypes.c_void_p(host_ptr), ctypes.c_size_t(n), 1,
                           ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    return f


print(f"pinned, one copy        : {timed(lambda: dst.copy_(src, non_blocking=True)):7.1f} us per 1.18 MB (back to back, includes the call)")
for k in (2, 4):
    print(f"pinned, {k} streams       : {timed(split(src, k)):7.1f} us")
print(f"pinned, raw cudaMemcpy  : {timed(raw(src.data_ptr())):7.1f} us")
print(f"write-combined, raw     : {timed(raw(p.value)):7.1f} us")
big = torch.empty(64 << 20, dtype=torch.uint8).pin_memory()
dbig = torch.empty(64 << 20, dtype=torch.uint8, device=dev)
t = timed(lambda: dbig.copy_(big, non_blocking=True), 20)
print(f"64 MB pinned copy       : {t:7.1f} us = {(64 << 20) / t / 1e3:.1f} GB/s")
