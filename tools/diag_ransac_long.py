import numpy as np, os, sys
sys.path.insert(0, "/root/repo")
import calib_b200
from calib_b200 import calib_ransac as cr
g = np.load("tests/golden/ransac_c4_long.npz")
prob = cr.RansacProblem(g["vc"], g["n"])
s, Rt = prob.fit_score(g["sn"].astype(np.int32), g["sm"].astype(np.int32), want_models=True)
Rt = Rt.cpu().numpy(); s = s.cpu().numpy()
dt = np.max(np.abs(Rt[:, :, :, 3:] - g["t_hyp"]), axis=(1, 2, 3)); dR = np.max(np.abs(Rt[:, :, :, :3] - g["R_hyp"]), axis=(1, 2, 3))
tm = np.max(np.abs(g["t_hyp"]), axis=(1, 2, 3))
rel = np.abs(s - g["scores"]) / g["scores"]
idx = np.argsort(-dt)[:8]
for h in idx: print(h, "dt %.2e dR %.2e |t|max %.2e score/best %.1f rel %.2e" % (dt[h], dR[h], tm[h], g["scores"][h] / g["scores"].min(), rel[h]))
print(prob.fit_stats())
