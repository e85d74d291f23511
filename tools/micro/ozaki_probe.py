"""Go / no-go probe for an Ozaki-split FP64 SYRK on the int8 tensor cores (stretch).

The SYRK of the LM step is S = Y Y^T with Y (397 x 637500) FP64: 1.007e11 algorithmic flops, 3.39 ms on the FP64 DMMA pipe.
An Ozaki scheme splits every (row-scaled) FP64 entry into p 7-bit slices held as int8 and forms S from the slice-pair
products with s + t < p: p (p + 1) / 2 int8 GEMMs of the same shape, each accumulated in int32 over K-chunks short enough
not to overflow (127^2 * 65536 < 2^31), recombined in FP64.  For the damped step to stay within 1e-8 the reduced system
needs ~43 bits -> p = 7 -> 28 slice-pair GEMMs (14-16 if the symmetric half is skipped).

This probe measures the best case for the GEMM part with the vendor library (torch._int_mm -> cuBLASLt int8, int32
accumulate) at the real shape, plus the FP64->int8 split pass as a plain torch expression of its memory traffic.  It is a
feasibility measurement only — nothing here is on the product path."""
import json, sys, time
import torch

def t_ms(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

dev = torch.device("cuda")
n, K = 400, 637504                      # 397 rows padded to 400, K padded to a multiple of 64
out = {"n": n, "K": K}
A = torch.randint(-127, 128, (n, K), dtype=torch.int8, device=dev)
Bt = A.t().contiguous()                 # (K, n)
for kc in (65536, 131072, K):
    kc = min(kc, K)
    chunks = [(a, min(a + kc, K)) for a in range(0, K, kc)]
    def run():
        acc = None
        for a, b in chunks:
            c = torch._int_mm(A[:, a:b].contiguous() if (b - a) != K else A, Bt[a:b])
            acc = c if acc is None else acc + c
        return acc
    # contiguous() of the chunk is a copy we would not pay in a real kernel: time the GEMMs alone on pre-cut chunks
    cutA = [A[:, a:b].contiguous() for a, b in chunks]; cutB = [Bt[a:b].contiguous() for a, b in chunks]
    def run2():
        for x, y in zip(cutA, cutB):
            torch._int_mm(x, y)
    ms = t_ms(run2)
    out[f"int8_gemm_ms_kchunk_{kc}"] = ms
    out[f"int8_tops_kchunk_{kc}"] = 2.0 * n * n * K / (ms * 1e-3) / 1e12
# all slice pairs at once: the 7 slice planes stacked into one (7 x 400) x K operand; a hand-written kernel would skip the
# blocks with s + t >= 7 (21 of 49), so its useful share of this GEMM is 28 / 49
As = torch.randint(-127, 128, (7 * n, K), dtype=torch.int8, device=dev)
Bs = As.t().contiguous()
ms_st = t_ms(lambda: torch._int_mm(As, Bs), 3)
out["stacked_2800x2800_gemm_ms"] = ms_st
out["stacked_tops"] = 2.0 * (7 * n) ** 2 * K / (ms_st * 1e-3) / 1e12
out["stacked_useful_28_of_49_ms"] = ms_st * 28 / 49
del As, Bs
# the 16 useful slice pairs (s >= t, s + t <= 6) as four thin library GEMMs per overflow-safe K-chunk:
#   t = 0: planes 0..6 (2800 rows) x plane 0,  t = 1: planes 1..5 x plane 1,  t = 2: planes 2..4 x plane 2,  t = 3: plane 3 x plane 3
As = torch.randint(-127, 128, (7 * n, K), dtype=torch.int8, device=dev)
kc = 131072
chunks = [(a, min(a + kc, K)) for a in range(0, K, kc)]
ops = []
for a, b in chunks:
    for t in range(4):
        rows = As[t * n:(7 - t) * n, a:b].contiguous()
        col = As[t * n:(t + 1) * n, a:b].t().contiguous()
        ops.append((rows, col))
def thin():
    for r, c in ops:
        torch._int_mm(r, c)
ms_thin = t_ms(thin, 3)
out["thin_16_pairs_20_gemms_ms"] = ms_thin
del As, ops
Y = torch.randn(397, 637500, dtype=torch.float64, device=dev)
def split():
    # traffic of the split pass: read Y once, write 7 int8 planes (row scale, truncate, subtract, repeat)
    r = Y
    for s in range(7):
        q = torch.trunc(r * 64.0)
        q.to(torch.int8)
        r = (r * 64.0 - q) if s < 6 else r
ms_split = t_ms(split, 2)
out["split_pass_torch_ms"] = ms_split
out["split_pass_ideal_ms_at_6.4TBs"] = (Y.numel() * 8 + 7 * Y.numel()) / 6.4e12 * 1e3
best = min(v for k, v in out.items() if k.startswith("int8_gemm_ms"))
out["slice_pair_gemms_p7"] = 28
out["gemm_total_ms_28"] = 28 * best
out["gemm_total_ms_16_symmetric"] = 16 * best
out["dmma_syrk_ms_today"] = 3.39
best_total = min(16 * best, out["stacked_useful_28_of_49_ms"], out["thin_16_pairs_20_gemms_ms"]) + out["split_pass_ideal_ms_at_6.4TBs"]
out["best_case_total_ms"] = best_total
out["verdict"] = ("no-go" if best_total > 3.39 else "worth a kernel")
print(json.dumps(out))
