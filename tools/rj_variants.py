#!/usr/bin/env python
"""Time the k_rj variants (env BAX_RJ_VARIANT) on one workload and check they agree bit for bit.

    python tools/rj_variants.py [--workload cfg4shard|cfg3|cfg1] [--variants 0,1,2,3] [--reps 20]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg4shard")
    ap.add_argument("--variants", default="0,1,2,3")
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    import torch
    import calib_b200  # noqa: F401
    from calib_b200 import ba
    sc, (sp2d, ss2d, sp3d, ss3d), w = bench.make_workload(a.workload, 1000)
    peak, _ = bench.measured_peaks()
    ref = None
    for v in [int(t) for t in a.variants.split(",")]:
        os.environ["BAX_RJ_VARIANT"] = str(v)
        prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, *w["lam"])
        if ref is None:
            x0 = prob.triangulate(sc.R_init, sc.t_init)
            th_np = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
        th = prob._dev_theta(th_np)
        out = prob.residual_jacobian(th, with_index=True)
        for _ in range(3):
            prob.residual_jacobian(th, with_index=False, out=(out[0], out[1], out[2], None, None))
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.reps):
            prob.residual_jacobian(th, with_index=False, out=(out[0], out[1], out[2], None, None))
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.reps
        cbar = prob.n_vis / prob.P
        gbs = (184.0 + 24.0 / cbar) * prob.n_vis / (ms * 1e-3) / 1e9
        same = None
        if ref is None:
            ref = [t.clone() for t in out]
        else:
            same = all(torch.equal(r, t) for r, t in zip(ref, out))
        print(json.dumps({"variant": v, "workload": a.workload, "ms": round(ms, 5), "GBps": round(gbs, 1),
                          "frac": round(gbs / peak, 4), "bit_identical_to_first": same, "n_vis": prob.n_vis}))
        del out
        prob.close()


if __name__ == "__main__":
    main()
