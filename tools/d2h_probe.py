"""Device-to-host copy of one step's result slab (16.6 MB): same page-locked buffer vs a ring, cudaHostAlloc vs
transparent-huge-page memory registered with cudaHostRegister (run on the GPU box)."""
import ctypes, mmap, time
import numpy as np, torch
NB = 16646144 + 256
src = torch.empty(NB, dtype=torch.uint8, device="cuda")
def tm(f, reps=40):
    for _ in range(5): f()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e3
def ring_test(bufs):
    k = [0]
    def f():
        k[0] += 1; bufs[k[0] % len(bufs)].copy_(src, non_blocking=True); torch.cuda.synchronize()
    return tm(f)
pinned = [torch.empty(NB, dtype=torch.uint8).pin_memory() for _ in range(3)]
print("cudaHostAlloc  same buffer : %.3f ms" % ring_test(pinned[:1]))
print("cudaHostAlloc  ring of 2   : %.3f ms" % ring_test(pinned[:2]))
print("cudaHostAlloc  ring of 3   : %.3f ms" % ring_test(pinned))
libc = ctypes.CDLL("libc.so.6", use_errno=True)
def thp_pinned(nbytes):
    sz = (nbytes + (2 << 20) - 1) // (2 << 20) * (2 << 20)
    m = mmap.mmap(-1, sz + (2 << 20), flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
    a = np.frombuffer(m, dtype=np.uint8)
    off = (-a.ctypes.data) % (2 << 20)
    a = a[off:off + sz]
    rc = libc.madvise(ctypes.c_void_p(a.ctypes.data), ctypes.c_size_t(sz), 14)  # MADV_HUGEPAGE
    a[:] = 0
    t = torch.from_numpy(a)
    r = torch.cuda.cudart().cudaHostRegister(a.ctypes.data, sz, 0)
    return t, m, rc, r
thp = [thp_pinned(NB) for _ in range(3)]
print("madvise rc", [x[2] for x in thp], "register", [str(x[3]) for x in thp], "is_pinned", [x[0].is_pinned() for x in thp])
try:
    print("AnonHugePages:", [l for l in open("/proc/meminfo") if "AnonHuge" in l][0].strip(), "| thp:", open("/sys/kernel/mm/transparent_hugepage/enabled").read().strip())
except Exception as ex: print(ex)
tb = [x[0][:NB] for x in thp]
print("THP+register   same buffer : %.3f ms" % ring_test(tb[:1]))
print("THP+register   ring of 3   : %.3f ms" % ring_test(tb))
# the CPU reading the buffer after the copy (what a learner does): sum of the first 4 MB
def ring_read(bufs):
    k = [0]
    def f():
        k[0] += 1; b = bufs[k[0] % len(bufs)]; b.copy_(src, non_blocking=True); torch.cuda.synchronize(); b[:1 << 22].numpy().sum()
    return tm(f)
print("cudaHostAlloc  ring of 3 + CPU read 4 MB : %.3f ms" % ring_read(pinned))
print("chunked ring of 3 (4 copies of 4 MB)     : ", end="")
k = [0]
def chunked():
    k[0] += 1; b = pinned[k[0] % 3]
    for c in range(4):
        lo = c * (NB // 4); hi = NB if c == 3 else (c + 1) * (NB // 4)
        b[lo:hi].copy_(src[lo:hi], non_blocking=True)
    torch.cuda.synchronize()
print("%.3f ms" % tm(chunked))
