#!/usr/bin/env python3
"""Host -> device copy rate for the e2e leg's 331 MB per step: default pinned memory (what torch.pin_memory gives), write-combined
pinned memory, one copy vs two streams, and the NUMA node of the buffer.  Prints GB/s per variant."""
import ctypes as C
import os
import subprocess
import sys

import torch

rt = C.CDLL("libcudart.so.12")
rt.cudaHostAlloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t, C.c_uint]
rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
NBYTES = 331_503_072
dev = torch.device("cuda", 0)
dst = torch.empty(NBYTES, dtype=torch.uint8, device=dev)
dst2 = torch.empty(NBYTES, dtype=torch.uint8, device=dev)


def host(flags):
    p = C.c_void_p()
    assert rt.cudaHostAlloc(C.byref(p), NBYTES, flags) == 0
    C.memset(p, 1, NBYTES)
    return p


def rate(fn, reps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return NBYTES * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9


s0 = torch.cuda.current_stream().cuda_stream
side = torch.cuda.Stream()
out = {}
for name, flags in (("default", 0), ("portable", 1), ("write_combined", 4)):
    p = host(flags)
    out[name + "_one_copy"] = rate(lambda: rt.cudaMemcpyAsync(dst.data_ptr(), p, NBYTES, 1, s0))
    third = (NBYTES // 3) & ~255

    def three():
        for k in range(3):
            rt.cudaMemcpyAsync(dst.data_ptr() + k * third, p.value + k * third, third, 1, s0)
    out[name + "_three_copies"] = rate(three)
    half = (NBYTES // 2) & ~255

    def two_streams():
        ev = torch.cuda.Event(); ev.record()
        side.wait_event(ev)
        rt.cudaMemcpyAsync(dst.data_ptr(), p, half, 1, s0)
        rt.cudaMemcpyAsync(dst.data_ptr() + half, p.value + half, NBYTES - half, 1, side.cuda_stream)
        ev2 = torch.cuda.Event(); ev2.record(side); torch.cuda.current_stream().wait_event(ev2)
    out[name + "_two_streams"] = rate(two_streams)
t = torch.empty(NBYTES, dtype=torch.uint8).pin_memory()
out["torch_pin_memory"] = rate(lambda: dst.copy_(t, non_blocking=True))
print({k: round(v, 2) for k, v in out.items()})
try:
    print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout[:1500])
    print(subprocess.run(["numactl", "-H"], capture_output=True, text=True, timeout=20).stdout[:800])
except Exception as e:
    print("topology query failed:", e)
