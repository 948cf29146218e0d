"""H2D copy rate of 24 MB from different kinds of pinned host memory (cudart through ctypes):
default pinned, write-combined, and the same split over two streams."""
import ctypes as C
import time

import torch

torch.cuda.init()
rt = C.CDLL("libcudart.so.12") if True else None
N = 24_000_128
dev = torch.empty(N, dtype=torch.uint8, device="cuda")
dptr = C.c_void_p(dev.data_ptr())


def alloc(flags):
    p = C.c_void_p()
    rc = rt.cudaHostAlloc(C.byref(p), C.c_size_t(N), C.c_uint(flags))
    assert rc == 0, rc
    C.memset(p, 1, N)
    return p


def timeit(fn, n=30):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for i in range(n + 5):
        torch.cuda.synchronize()
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        if i >= 5:
            ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


st0 = torch.cuda.current_stream().cuda_stream
s2 = torch.cuda.Stream()
for name, flags in [("default pinned", 0), ("portable", 1), ("write-combined", 4)]:
    h = alloc(flags)

    def one():
        rt.cudaMemcpyAsync(dptr, h, C.c_size_t(N), C.c_int(1), C.c_void_p(st0))

    ms = timeit(one)
    print(f"{name:16s} one copy      {ms:.4f} ms  {N / ms / 1e6:.1f} GB/s")
    half = N // 2

    def two():
        rt.cudaMemcpyAsync(dptr, h, C.c_size_t(half), C.c_int(1), C.c_void_p(st0))
        with torch.cuda.stream(s2):
            s2.wait_stream(torch.cuda.current_stream())
        rt.cudaMemcpyAsync(C.c_void_p(dptr.value + half), C.c_void_p(h.value + half), C.c_size_t(half),
                           C.c_int(1), C.c_void_p(s2.cuda_stream))
        torch.cuda.current_stream().wait_stream(s2)

    ms = timeit(two)
    print(f"{name:16s} two streams   {ms:.4f} ms  {N / ms / 1e6:.1f} GB/s")
t = torch.empty(N, dtype=torch.uint8).pin_memory()
ms = timeit(lambda: dev.copy_(t, non_blocking=True))
print(f"torch pin_memory one copy      {ms:.4f} ms  {N / ms / 1e6:.1f} GB/s")
