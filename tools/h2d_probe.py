"""Host->device copy bandwidth on this box, the ceiling of the end-to-end figure (1.536 MB per 640x480 RGB-D frame):
one stream, several streams, chunked copies, pageable memory, write-combined pinned memory, and NUMA placement."""
import ctypes, json, os, sys, time
import numpy as np, torch

dev = torch.device("cuda", 0)
MB = 1 << 20
N = 96 * MB
out = {}

def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps

dst = torch.empty(N, dtype=torch.uint8, device=dev)
pin = torch.empty(N, dtype=torch.uint8).pin_memory()
pin.fill_(7)
out["pinned_1stream_GBs"] = N / timed(lambda: dst.copy_(pin, non_blocking=True)) / 1e9
pageable = torch.empty(N, dtype=torch.uint8); pageable.fill_(3)
out["pageable_GBs"] = N / timed(lambda: dst.copy_(pageable)) / 1e9
for k in (2, 4, 8):
    streams = [torch.cuda.Stream() for _ in range(k)]
    c = N // k
    def multi():
        for i, s in enumerate(streams):
            with torch.cuda.stream(s):
                dst[i * c:(i + 1) * c].copy_(pin[i * c:(i + 1) * c], non_blocking=True)
    out[f"pinned_{k}streams_GBs"] = N / timed(multi) / 1e9
def chunked():
    c = 4 * MB
    for i in range(0, N, c):
        dst[i:i + c].copy_(pin[i:i + c], non_blocking=True)
out["pinned_1stream_4MB_chunks_GBs"] = N / timed(chunked) / 1e9
# D2H
back = torch.empty(N, dtype=torch.uint8).pin_memory()
out["d2h_pinned_GBs"] = N / timed(lambda: back.copy_(dst, non_blocking=True)) / 1e9
# write-combined / portable pinned memory through the runtime
rt = torch.cuda.cudart()
try:
    libcudart = ctypes.CDLL("libcudart.so.12")
    for name, flag in (("wc", 4), ("portable", 1), ("mapped", 2)):
        p = ctypes.c_void_p()
        rc = libcudart.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(N), ctypes.c_uint(flag))
        if rc != 0:
            out[f"hostalloc_{name}"] = f"error {rc}"; continue
        ctypes.memset(p, 5, N)
        def cp():
            libcudart.cudaMemcpyAsync(ctypes.c_void_p(dst.data_ptr()), p, ctypes.c_size_t(N), ctypes.c_int(1), ctypes.c_void_p(0))
        out[f"hostalloc_{name}_GBs"] = N / timed(cp) / 1e9
        libcudart.cudaFreeHost(p)
except Exception as e:
    out["hostalloc_error"] = repr(e)
# a kernel reading pinned host memory directly (zero-copy) through a mapped tensor
try:
    hp = torch.empty(N // 4, dtype=torch.int32).pin_memory()
    out["note_zero_copy"] = "not probed (torch cannot view pinned host memory as a device tensor)"
except Exception as e:
    out["zero_copy_error"] = repr(e)
out["cpus"] = os.cpu_count()
try:
    out["affinity"] = sorted(os.sched_getaffinity(0))[:4] + ["..."] + [len(os.sched_getaffinity(0))]
    out["numa_nodes"] = sorted(d for d in os.listdir("/sys/devices/system/node") if d.startswith("node"))
    out["gpu_numa"] = open([p for p in [f"/sys/bus/pci/devices/{torch.cuda.get_device_properties(0).pci_bus_id if hasattr(torch.cuda.get_device_properties(0),'pci_bus_id') else ''}/numa_node"] if os.path.exists(p)][0]).read().strip() if False else None
except Exception as e:
    out["numa_error"] = repr(e)
os.system("nvidia-smi topo -m > /tmp/topo.txt 2>&1; nvidia-smi -q | grep -i -A4 'GPU Link Info' | head -20 >> /tmp/topo.txt 2>&1")
out["topo"] = open("/tmp/topo.txt").read()[:3000]
print(json.dumps(out, indent=1))
