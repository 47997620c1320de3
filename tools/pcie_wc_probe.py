"""H2D bandwidth from default page-locked memory against write-combined page-locked memory (cudaHostAllocWriteCombined),
one large copy and 32-frame chunks (what orbx_extract_batch issues), CUDA-event timed."""
import ctypes as C
import torch

rt = C.CDLL("libcudart.so.12")
rt.cudaHostAlloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t, C.c_uint]
rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
torch.cuda.init()
dev = torch.device("cuda", 0)
n = 512 * 640 * 480
d = torch.empty(n, dtype=torch.uint8, device=dev)
for name, flags in (("default", 0), ("write-combined", 4), ("default", 0), ("write-combined", 4)):
    p = C.c_void_p()
    assert rt.cudaHostAlloc(C.byref(p), n, flags) == 0
    C.memset(p, 7, n)
    for label, chunk in (("one copy", n), ("32-frame chunks", 32 * 640 * 480)):
        s = torch.cuda.current_stream().cuda_stream
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for it in range(4):
                for off in range(0, n, chunk):
                    rt.cudaMemcpyAsync(C.c_void_p(d.data_ptr() + off), C.c_void_p(p.value + off), min(chunk, n - off), 1, C.c_void_p(s))
            e1.record()
            torch.cuda.synchronize()
        print("%-15s %-16s %.1f GB/s" % (name, label, 4 * n / e0.elapsed_time(e1) / 1e6), flush=True)
    rt.cudaFreeHost(p)
