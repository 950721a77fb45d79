"""What a first-time device allocation costs on this box: cudaMalloc / cudaFree of growing sizes, in a process where torch's
caching allocator already holds memory (like the bench) and in one where it does not.  usage: python tools/malloc_cost.py [hold_gb]"""
import ctypes as C, sys, time
import torch
hold = float(sys.argv[1]) if len(sys.argv) > 1 else 0.0
torch.cuda.init(); torch.zeros(1, device="cuda")
keep = torch.empty(int(hold * 2**30), dtype=torch.uint8, device="cuda") if hold > 0 else None
rt = C.CDLL("libcudart.so.12")
rt.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
rt.cudaFree.argtypes = [C.c_void_p]
print(f"torch holds {hold} GB")
for mb in (1, 16, 64, 256, 1024, 2048, 4096):
    ts = []
    for rep in range(3):
        p = C.c_void_p()
        t0 = time.perf_counter(); rc = rt.cudaMalloc(C.byref(p), mb << 20); t1 = time.perf_counter()
        rt.cudaFree(p); t2 = time.perf_counter()
        ts.append((1e3 * (t1 - t0), 1e3 * (t2 - t1)))
    print(f"  {mb:5d} MB: cudaMalloc {ts[0][0]:8.3f} / {ts[1][0]:8.3f} / {ts[2][0]:8.3f} ms   cudaFree {ts[0][1]:8.3f} / {ts[2][1]:8.3f} ms  rc={rc}")
# many medium blocks, as the hierarchy build makes them
t0 = time.perf_counter(); ps = []
for i in range(100):
    p = C.c_void_p(); rt.cudaMalloc(C.byref(p), 64 << 20); ps.append(p)
t1 = time.perf_counter()
for p in ps: rt.cudaFree(p)
print(f"  100 x 64 MB: {1e3 * (t1 - t0):.1f} ms to allocate, {1e3 * (time.perf_counter() - t1):.1f} ms to free")
