import sys, time
sys.path.insert(0, "/root/repo")
import bench, torch
import calib_b200
from calib_b200 import ba
sc, (sp2d, ss2d, sp3d, ss3d), w = bench.make_workload("cfg4shard", 1000)
C, N, J = ss2d.shape
y_h = torch.from_numpy(sp2d.reshape(C, N * J, 2)).pin_memory()
s_h = torch.from_numpy(ss2d.reshape(C, N * J)).pin_memory()
prob = ba.BAProblem(sc.K, y_h, s_h.reshape(C, N, J), sp3d, ss3d, ba.OP_BONE, *w["lam"])
x0 = prob.triangulate(sc.R_init, sc.t_init)
th = prob._dev_theta(ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3)))
def T(f, n=5):
    f(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
y2, s2 = torch.empty_like(prob.y), torch.empty_like(prob.s)
def up(): y2.copy_(y_h, non_blocking=True); s2.copy_(s_h, non_blocking=True)
print("upload alone ms", T(up))
def comp():
    s = ba.TRFSolver(prob, th); s.iterate(relinearize=False)
print("compute alone ms", T(comp))
cs = torch.cuda.Stream()
def both():
    with torch.cuda.stream(cs): up()
    comp()
print("both ms", T(both))
def lin(): prob.linearize(th)
print("linearize ms", T(lin))
sol = ba.TRFSolver(prob, th)
def it(): 
    global sol
    sol.iterate()
    if sol.done(): sol = ba.TRFSolver(prob, th)
print("iterate ms", T(it, 10))
