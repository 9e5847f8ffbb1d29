import ctypes, torch, time
rt = ctypes.CDLL("libcudart.so.12")
n = 67 * 1024 * 1024
dev = torch.empty(n, dtype=torch.uint8, device="cuda")
dev2 = torch.empty(n, dtype=torch.uint8, device="cuda")
def bw(src_ptr, label, two=False):
    s = torch.cuda.Stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for rep in range(2):
        with torch.cuda.stream(s):
            e0.record(s)
            for i in range(20):
                rt.cudaMemcpyAsync(ctypes.c_void_p(dev.data_ptr()), ctypes.c_void_p(src_ptr), ctypes.c_size_t(n), 1, ctypes.c_void_p(s.cuda_stream))
                rt.cudaMemcpyAsync(ctypes.c_void_p(dev2.data_ptr()), ctypes.c_void_p(src_ptr), ctypes.c_size_t(n), 1, ctypes.c_void_p(s.cuda_stream))
            e1.record(s)
        torch.cuda.synchronize()
    print(label, "%.1f GB/s" % (40 * n / (e0.elapsed_time(e1) * 1e-3) / 1e9))
t = torch.empty(n, dtype=torch.uint8).pin_memory(); t.fill_(1)
bw(t.data_ptr(), "torch pin_memory")
for flags, name in ((0, "cudaHostAlloc default"), (4, "write-combined"), (1, "portable")):
    p = ctypes.c_void_p()
    rc = rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(n), flags)
    assert rc == 0, rc
    ctypes.memset(p, 1, n)
    bw(p.value, name)
    rt.cudaFreeHost(p)
# one big copy of 134 MB vs two of 67
