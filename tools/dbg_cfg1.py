import sys, os
sys.path.insert(0, "/root/repo")
import bench, torch, math
import calib_b200
from calib_b200 import ba
sc, (sp2d, ss2d, sp3d, ss3d), w = bench.make_workload("cfg1", 1000)
prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, *w["lam"])
x0 = prob.triangulate(sc.R_init, sc.t_init)
th = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
orig = prob.trial
def trial(alpha, need=True):
    r = orig(alpha, need)
    print("  trial alpha=%.6e need=%s -> pn2=%.6e gp=%.6e ptq=%.6e" % (alpha, need, *r))
    return r
prob.trial = trial
sol = ba.TRFSolver(prob, th, ftol=0.0, xtol=0.0, gtol=0.0)
for it in range(16):
    print("iter", it, "cost", sol.cost, "Delta", sol.Delta, "alpha", sol.alpha, "g2", sol.g_2, "hscale", prob.h_scale)
    try:
        sol.iterate()
    except ZeroDivisionError as e:
        print("ZERO DIV"); break
