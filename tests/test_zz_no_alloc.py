"""rsx_argmin_static (include/rsx.h): the winner search without any allocation, its scratch in a static device array.
No counterpart in the reference; the check is the existing entry point
rsx_argmin on the same scores — index and score are integer / copy work, so the bar is bit-exact.

Collected last on purpose."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import calib_b200  # noqa: F401
from calib_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHUNK, MAX_H, SCRATCH = 4096, 1 << 24, 65792      # RSX_STATIC_* of include/rsx.h


def _header_defines():
    text = open(os.path.join(ROOT, "include", "rsx.h")).read()
    out = {}
    for name in ("RSX_STATIC_CHUNK", "RSX_STATIC_MAX_H", "RSX_STATIC_SCRATCH_BYTES"):
        out[name] = eval(re.search(rf"#define\s+{name}\s+(\([^)]*\)|\S+)", text).group(1))
    return out


def test_limits_are_the_declared_ones_and_oversize_is_rejected():
    d = _header_defines()
    assert (d["RSX_STATIC_CHUNK"], d["RSX_STATIC_MAX_H"], d["RSX_STATIC_SCRATCH_BYTES"]) == (CHUNK, MAX_H, SCRATCH)
    assert SCRATCH == 256 + 16 * (MAX_H // CHUNK)
    L = _lib.lib()
    # argument checks come before any CUDA call: no GPU needed (the pointers are never dereferenced)
    assert L.rsx_argmin_static(0, 256, MAX_H + 1, 256, None) == -1
    msg = L.bax_last_error().decode()
    assert "RSX_STATIC_MAX_H" in msg and str(MAX_H + 1) in msg
    assert L.rsx_argmin_static(0, 256, 0, 256, None) == -1
    assert L.rsx_argmin_static(0, None, 10, 256, None) == -1
    assert L.rsx_argmin_static(0, 256, 10, None, None) == -1


def test_source_calls_no_allocator():
    src = open(os.path.join(ROOT, "extrinsic-camera-calibration-from-a-moving-person_b200", "csrc", "rsx_static.cu")).read()
    code = re.sub(r"//[^\n]*", "", src)
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMallocAsync", "malloc(", "new ", "torch"):
        assert banned not in code, banned
    assert "cudaGetSymbolAddress" in code and "static __device__ __align__(256)" in code


def _both(L, h, score):
    """(index, score) from the existing entry point and from the static one, on the same device scores."""
    import torch
    dev = score.device
    b, s = C.c_int64(-7), C.c_double(-7.0)
    stream = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(L.rsx_argmin(h, score.data_ptr(), score.numel(), C.byref(b), C.byref(s), stream), "rsx_argmin")
    out = torch.full((2,), -7.0, dtype=torch.float64, device=dev)
    _lib.check(L.rsx_argmin_static(dev.index or 0, score.data_ptr(), score.numel(), out.data_ptr(), stream),
               "rsx_argmin_static")
    o = out.cpu().numpy()
    return (b.value, s.value), (int(o[0]), float(o[1]))


@pytest.mark.gpu
def test_static_argmin_matches_rsx_argmin_on_calls_in_a_row():
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    L = _lib.lib()
    h = C.c_void_p()
    _lib.check(L.rsx_create(0, C.byref(h)), "rsx_create")
    try:
        g = torch.Generator(device="cpu").manual_seed(20261105)
        # call 1: many CTAs, the minimum in the last chunk, duplicated values (the first one has to win)
        H1 = 37 * CHUNK + 1234
        s1 = torch.rand(H1, generator=g, dtype=torch.float64) + 1.0
        s1[H1 - 100] = 0.25; s1[H1 - 50] = 0.25; s1[5] = float("nan"); s1[6] = float("inf")
        ref, got = _both(L, h, s1.to(dev))
        assert ref == (H1 - 100, 0.25) and got == ref
        # call 2 right after it, FEWER CTAs: the scratch still holds call 1's partials (0.25 among them), which must
        # not leak into this answer
        H2 = 3 * CHUNK + 17
        s2 = torch.rand(H2, generator=g, dtype=torch.float64) + 2.0
        s2[CHUNK + 3] = 1.5
        ref, got = _both(L, h, s2.to(dev))
        assert ref == (CHUNK + 3, 1.5) and got == ref
        # call 3: a single score; call 4: nothing can win (NaN / +inf only) -> index -1, score +inf
        ref, got = _both(L, h, torch.tensor([3.0], dtype=torch.float64, device=dev))
        assert ref == (0, 3.0) and got == ref
        s4 = torch.full((2 * CHUNK,), float("inf"), dtype=torch.float64); s4[::3] = float("nan")
        ref, got = _both(L, h, s4.to(dev))
        assert ref[0] == -1 and got[0] == -1 and np.isinf(got[1]) and np.isinf(ref[1])
        # random sizes incl. exact chunk multiples, random data with ties from quantisation
        for H in (CHUNK, CHUNK + 1, 2 * CHUNK - 1, 100_000, 1_000_003):
            s = (torch.randint(0, 50_000, (H,), generator=g).to(torch.float64) / 7.0).to(dev)
            ref, got = _both(L, h, s)
            assert got == ref
            assert got[0] == int((s.cpu() == s.cpu().min()).nonzero()[0].item())   # first of the equal minima
    finally:
        L.rsx_destroy(h)


@pytest.mark.gpu
def test_static_argmin_at_the_maximum_size():
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    L = _lib.lib()
    h = C.c_void_p()
    _lib.check(L.rsx_create(0, C.byref(h)), "rsx_create")
    try:
        s = torch.full((MAX_H,), 9.0, dtype=torch.float64, device=dev)
        s[MAX_H - 1] = 8.0
        ref, got = _both(L, h, s)
        assert ref == (MAX_H - 1, 8.0) and got == ref
        s[MAX_H - 1] = 9.0                                   # all equal: the first index wins
        ref, got = _both(L, h, s)
        assert ref == (0, 9.0) and got == ref
        out = torch.zeros(2, dtype=torch.float64, device=dev)
        assert L.rsx_argmin_static(0, s.data_ptr(), MAX_H + 1, out.data_ptr(), None) == -1
    finally:
        L.rsx_destroy(h)
